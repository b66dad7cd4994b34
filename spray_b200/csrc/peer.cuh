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
// one thread: wait until every peer has signalled `target` barriers; ~4 s timeout instead of a hung GPU
static __global__ void p2p_wait_kernel(volatile unsigned long long* flag, unsigned long long target, unsigned int* err) {
  if (*(volatile unsigned int*)err) return;      // an earlier barrier already gave up: do not wait 4 s again
  const long long t0 = clock64();
  while (*flag < target) {
    if (clock64() - t0 > 8000000000ll) { *err = 1u; break; }
    __nanosleep(200);
  }
  __threadfence_system();
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
