// peer.cuh — device-side barrier and append helpers shared by the peer-memory in-situ tiles (insitu2.cu: dense wave
// arrays; insitu3.cu: ray forwarding).  Every rank maps every other rank's region (CUDA IPC over NVLink / NVSwitch,
// comm_shared_alloc in comm.cu); offset 0 of a region is its barrier counter.
#pragma once
#include <stdlib.h>

#include <string>
#include <utility>
#include <vector>

#include "internal.cuh"

namespace spraygpu {

struct Comm;
int comm_shared_alloc(Comm* c, size_t bytes, void** local, std::vector<void*>* mapped, int only_rank);
void comm_shared_free(Comm* c, void* local, std::vector<void*>* mapped);
int comm_host_barrier(Comm* c);
int comm_allreduce_u64_sum(Comm* c, unsigned long long* p, size_t n);
int comm_rank(const Comm* c);
int comm_size(const Comm* c);
Device* comm_device(const Comm* c);
void** comm_p2p_slot(Comm* c, void (***free_fn)(Comm*, void*), int which);

#define P2P_MAX_RANKS 16
#define P2P_FLAG_BYTES 256

struct PeerPtrs { char* base[P2P_MAX_RANKS]; };

// ---- device-side barrier ------------------------------------------------------------------------------------------
// Called by EVERY thread at the end of a kernel whose peer stores must be visible before the peers go on.
__device__ __forceinline__ void grid_signal_peers(unsigned int* done, const PeerPtrs& P, int nranks, int me) {
  __syncthreads();                         // the CTA's peer stores / atomics happen-before thread 0's fence ...
  if (threadIdx.x == 0) {
    __threadfence_system();                // ... which is cumulative: they are performed system-wide before the signal
    const unsigned int total = gridDim.x * gridDim.y;
    if (atomicAdd(done, 1u) == total - 1u) {
      *done = 0u;                          // ready for the next signalling kernel (stream order)
      __threadfence_system();
      for (int r = 0; r < nranks; ++r)
        if (r != me) atomicAdd_system((unsigned long long*)P.base[r], 1ull);
    }
  }
}
static __global__ void p2p_signal_kernel(PeerPtrs P, int nranks, int me) {
  __threadfence_system();
  for (int r = 0; r < nranks; ++r)
    if (r != me) atomicAdd_system((unsigned long long*)P.base[r], 1ull);
}
// one thread: wait until every peer has signalled `target` barriers; gives up after `timeout` SM cycles (default ~4 s,
// p2p_timeout_cycles) instead of hanging the GPU, and says so in bit 0 of the mapped error word
#define P2P_ERR_TIMEOUT 1u
#define P2P_ERR_OVERFLOW 2u
static __global__ void p2p_wait_kernel(volatile unsigned long long* flag, unsigned long long target, unsigned int* err, long long timeout) {
  if (*(volatile unsigned int*)err & P2P_ERR_TIMEOUT) return;      // an earlier barrier already gave up: do not wait again
  const long long t0 = clock64();
  while (*flag < target) {
    if (clock64() - t0 > timeout) { *err = *(volatile unsigned int*)err | P2P_ERR_TIMEOUT; break; }   // one thread, stream order: no atomic on mapped host memory
    __nanosleep(200);
  }
  __threadfence_system();
}
// SPRAY_PEER_TIMEOUT_S: seconds a rank waits for its peers at a device-side barrier (default 4).  Raise it when one
// rank is legitimately late: under ncu / compute-sanitizer (kernels serialised and replayed), a long host-side step
// between tiles on one rank, the first lazy module load.  Counted in SM cycles at a nominal 2 GHz.
static inline long long p2p_timeout_cycles() {
  static long long v = 0;
  if (!v) {
    const char* e = getenv("SPRAY_PEER_TIMEOUT_S");
    double sec = e ? atof(e) : 4.0;
    if (!(sec > 0.0)) sec = 4.0;
    if (sec > 3600.0) sec = 3600.0;
    v = (long long)(sec * 2.0e9);
  }
  return v;
}
// The mapped error word at the start of a tile call: report what the previous tile left there, ONCE, and clear it — the
// communicator stays usable (a rank that was merely late has signalled by now; the barrier counters are cumulative).
static inline int p2p_take_error(unsigned int* h_err, const char* who) {
  if (!h_err || *h_err == 0u) return SPRAYGPU_OK;
  const unsigned int e = *h_err;
  *h_err = 0u;
  if (e & P2P_ERR_TIMEOUT) {
    set_error("%s: a peer barrier of the previous tile timed out (a rank died, the ranks' call sequences differ, or a rank was more "
              "than SPRAY_PEER_TIMEOUT_S late); that tile's pixels are incomplete%s", who,
              (e & P2P_ERR_OVERFLOW) ? "; an inbox also overflowed" : "");
    return SPRAYGPU_ERR_STATE;
  }
  set_error("%s: an inbox overflowed in the previous tile (rays were dropped); use smaller blocking tiles", who);
  return SPRAYGPU_ERR_LIMIT;
}

// full-warp call: one atomic per warp, returns this lane's slot (valid where pred)
__device__ __forceinline__ uint32_t warp_append(unsigned int* counter, bool pred) {
  const uint32_t b = __ballot_sync(0xffffffffu, pred);
  if (!b) return 0u;
  const int lane = threadIdx.x & 31, leader = __ffs(b) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(counter, (unsigned int)__popc(b));
  base = __shfl_sync(0xffffffffu, base, leader);
  return base + __popc(b & ((1u << lane) - 1u));
}

// SPRAY_INSITU_TIMING=1: CUDA events between the phases of a tile, printed per rank at the end of the call (a
// development aid: it synchronises the stream, so timed benchmark runs leave it off)
struct PhaseTimer {
  bool on = false;
  cudaStream_t st = nullptr;
  std::vector<std::pair<const char*, cudaEvent_t>> ev;
  explicit PhaseTimer(cudaStream_t s) : st(s) { const char* e = getenv("SPRAY_INSITU_TIMING"); on = e && *e == '1'; mark("start"); }
  void mark(const char* name) {
    if (!on) return;
    cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st);
    ev.push_back(std::make_pair(name, e));
  }
  void report(const char* what, int rank) {
    if (!on) return;
    cudaStreamSynchronize(st);
    std::string line = std::string(what) + " rank " + std::to_string(rank) + ":";
    for (size_t i = 1; i < ev.size(); ++i) {
      float ms = 0.f; cudaEventElapsedTime(&ms, ev[i - 1].second, ev[i].second);
      char buf[64]; snprintf(buf, sizeof(buf), " %s %.0f", ev[i].first, ms * 1e3f);
      line += buf;
    }
    float tot = 0.f; cudaEventElapsedTime(&tot, ev.front().second, ev.back().second);
    fprintf(stderr, "%s | total %.0f us\n", line.c_str(), tot * 1e3f);
    for (auto& e : ev) cudaEventDestroy(e.second);
    ev.clear();
  }
};

}  // namespace spraygpu
