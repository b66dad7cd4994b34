// comm.cc — see spray/comm.h; also the pinned-memory helpers of render_types.h (cudaHostAlloc through
// the CUDA runtime that libspraygpu already links; plain aligned memory when no device is present so
// the CPU-only unit tests of the host logic still run).
#include "spray/comm.h"

#include <cuda_runtime_api.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <unordered_set>

#include "spray/check.h"
#include "spray/render_types.h"

namespace spray {

static std::unordered_set<void*>& pinnedSet() { static std::unordered_set<void*> s; return s; }

void* allocPinned(std::size_t bytes) {
  if (bytes == 0) bytes = 16;
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) == cudaSuccess && p) {
    pinnedSet().insert(p);
    return p;
  }
  cudaGetLastError();
  if (posix_memalign(&p, 64, bytes) != 0) return nullptr;
  return p;
}

void freePinned(void* p) {
  if (!p) return;
  auto it = pinnedSet().find(p);
  if (it != pinnedSet().end()) { pinnedSet().erase(it); cudaFreeHost(p); }
  else std::free(p);
}

namespace mpi {

static int g_rank = 0, g_size = 1, g_local = 0;

static int envInt(const char* name, int dflt) {
  const char* v = std::getenv(name);
  return v ? std::atoi(v) : dflt;
}

void init() {
  g_rank = envInt("RANK", 0);
  g_size = envInt("WORLD_SIZE", 1);
  g_local = envInt("LOCAL_RANK", g_rank);
  SPRAY_CHECK(g_size >= 1 && g_rank >= 0 && g_rank < g_size, "bad RANK/WORLD_SIZE");
}
int rank() { return g_rank; }
int size() { return g_size; }
int localRank() { return g_local; }

spraygpu_comm createComm(spraygpu_device dev) {
  if (g_size == 1) return nullptr;
  char id[SPRAYGPU_UNIQUE_ID_BYTES];
  char path[256];
  const char* port = std::getenv("MASTER_PORT");
  const char* run = std::getenv("TORCHELASTIC_RUN_ID");
  std::snprintf(path, sizeof(path), "/dev/shm/spraygpu_ncclid_%s_%s_%d", port ? port : "0", run ? run : "none",
                (int)getppid());
  if (g_rank == 0) {
    SPRAY_GPU_CALL(spraygpu_comm_unique_id(id));
    char tmp[300];
    std::snprintf(tmp, sizeof(tmp), "%s.tmp", path);
    FILE* f = std::fopen(tmp, "wb");
    SPRAY_CHECK(f != nullptr, tmp);
    std::fwrite(id, 1, sizeof(id), f);
    std::fclose(f);
    std::rename(tmp, path);  // atomic publish
  } else {
    bool ok = false;
    for (int tries = 0; tries < 6000 && !ok; ++tries) {  // up to 60 s
      FILE* f = std::fopen(path, "rb");
      if (f) {
        ok = std::fread(id, 1, sizeof(id), f) == sizeof(id);
        std::fclose(f);
      }
      if (!ok) std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    SPRAY_CHECK(ok, "rank " << g_rank << ": no NCCL id at " << path);
  }
  spraygpu_comm comm = nullptr;
  SPRAY_GPU_CALL(spraygpu_comm_create(dev, g_size, g_rank, id, &comm));
  if (g_rank == 0) std::remove(path);  // every rank has joined once comm_create returns
  return comm;
}

}  // namespace mpi
}  // namespace spray
