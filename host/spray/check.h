// check.h — the reference never propagates errors: CHECK*/LOG(FATAL) abort the process
// (glog; e.g. src/render/trimesh_buffer.cc:68-113, src/io/ply_loader.cc:170-176).  The host mirror keeps
// that behaviour.  For the CPU-only unit tests the abort can be turned into a C++ exception
// (spray::setThrowOnFatal) so that a test can observe "the reference would have died here".
#pragma once
#include <cstdio>
#include <cstdlib>
#include <sstream>
#include <stdexcept>
#include <string>

namespace spray {

inline bool& throwOnFatalFlag() { static bool f = false; return f; }
inline void setThrowOnFatal(bool on) { throwOnFatalFlag() = on; }

struct FatalError : std::runtime_error { using std::runtime_error::runtime_error; };

[[noreturn]] inline void fatal(const char* file, int line, const std::string& msg) {
  std::ostringstream os;
  os << "[FATAL] " << file << ":" << line << ": " << msg;
  if (throwOnFatalFlag()) throw FatalError(os.str());
  std::fprintf(stderr, "%s\n", os.str().c_str());
  std::abort();
}

}  // namespace spray

#define SPRAY_CHECK(cond, msg)                                                   \
  do {                                                                           \
    if (!(cond)) {                                                               \
      std::ostringstream os__;                                                   \
      os__ << "Check failed: " #cond " " << msg;                                 \
      ::spray::fatal(__FILE__, __LINE__, os__.str());                            \
    }                                                                            \
  } while (0)

#define SPRAY_FATAL(msg)                              \
  do {                                                \
    std::ostringstream os__;                          \
    os__ << msg;                                      \
    ::spray::fatal(__FILE__, __LINE__, os__.str());   \
  } while (0)

// C-ABI calls: abort on a non-zero return code, printing spraygpu_last_error()
#define SPRAY_GPU_CALL(call)                                                                     \
  do {                                                                                           \
    int rc__ = (call);                                                                           \
    if (rc__ != 0) SPRAY_FATAL(#call " failed (" << rc__ << "): " << spraygpu_last_error());     \
  } while (0)
