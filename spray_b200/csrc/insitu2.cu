// insitu2.cu — the in-situ (domain-partitioned) blocking tile over PEER MEMORY: spraygpu_trace_tile_insitu_p2p.
//
// Reference: insitu::{SingleThread,MultiThread}Tracer::trace (src/insitu/insitu_singlethread_tracer.inl:381-466): per
// ray depth one super-step — process rays in the local domains, MPI_Allreduce(MIN) of the t-buffer
// (insitu_vbuf.h:109-117), MPI_Allreduce(BOR) of the occlusion buffer (:119-129), retire shadow rays, keep the
// secondaries whose parent t survived.  insitu.cu restates that with five dense NCCL all-reduces per wave
// (8 + 12 + 1 + 24 B per SAMPLE, hit or not) and lands at 0.26x of ONE GPU at 8 GPUs.
//
// Here every rank maps every other rank's wave buffers (CUDA IPC over NVLink / NVSwitch) and the exchange happens
// inside the compute kernels, only for samples that HIT:
//   K1 trace   : each rank traces the wave against ITS domains; a hit does a 64-bit atomicMin of the packed
//                (t | domain | prim) key straight into EVERY rank's key array (replaces the u64 min-all-reduce)
//   K2 capture : the rank that owns the winning domain stores the hit's (colour, shading normal) — 16 B — into every
//                rank's attribute array (replaces the 12 + 24 B sum-all-reduces of shadow direction / bounce record)
//   K3 shade   : EVERY rank shades every hit from (key, attributes) — identical arithmetic, so shadow rays, bounce rays
//                and radiance are the same everywhere without being sent — traces the shadow rays against ITS domains
//                and stores a 1 into the HOME rank's occlusion byte when occluded (replaces the u8 max-all-reduce)
//   K4 retire  : the home rank of a sample (interleaved blocks of image rows) adds the unoccluded radiance into ITS
//                frame buffer rows; spraygpu_image_flush pushes those rows to the root (image.cu)
// Between the kernels the ranks meet at device-side barriers: the last CTA of a producing kernel adds 1 to a counter in
// every peer's region (system-scope atomics after a system fence); a one-thread kernel spins on the local counter.
// One rank per GPU only (B200_PROFILING.md: kernels that wait on each other must not share a GPU).
#include <vector>

#include "internal.cuh"
#include "traverse.cuh"
#include "shade.cuh"
#include "peer.cuh"

namespace spraygpu {

struct P2PState {
  size_t cap_n = 0;            // samples the buffers hold
  int cap_passes = 0;          // light-sample passes per hit they hold
  size_t region_bytes = 0;
  void* local = nullptr;       // the peer-mapped region of THIS rank
  std::vector<void*> mapped;   // [nranks]
  size_t o_keys[2] = {0, 0}, o_cap = 0, o_occ = 0;
  unsigned long long seq = 0;  // barriers passed so far (same on every rank)
  // rank-local buffers
  spraygpu_ray* rays[2] = {nullptr, nullptr};
  float* w[2] = {nullptr, nullptr};
  uint32_t* list[2] = {nullptr, nullptr};   // samples of the next wave (those that bounced)
  uint4* ret = nullptr;        // retire list: (pass * n + sample, radiance) of the home samples' shadow rays
  unsigned int* d_ctl = nullptr;           // CTL_* words: "CTAs finished" counter, work counters, list lengths
  int grid_trace = 0, grid_shade = 0;      // persistent grids (occupancy x SMs), queried once per shared-memory size
  size_t grid_smem = (size_t)-1;
  unsigned int* h_err = nullptr;           // mapped pinned: set when a barrier wait timed out
  unsigned int* h_err_dev = nullptr;
};

}  // namespace spraygpu

namespace spraygpu {

// control words of one rank (device memory, rank-local)
enum { CTL_DONE = 0, CTL_WORK1, CTL_WORK3, CTL_RET, CTL_LIVE0, CTL_LIVE1, CTL_WORDS = 8 };
#define P2P_QCAP 64              // per-warp compaction queue entries (pop 32, refill below 32 -> <= 63)
#define P2P_THREADS 128
#define P2P_MINBLOCKS 4

struct P2PWave {
  const DomainEntry* table;
  int nd;
  const TlasNode* tlas;
  int tlas_n;
  spraygpu_shade shade;
  spraygpu_tile tile;
  int image_w, spp;
  int depth;               // depth of the wave being traced / shaded (eye rays: 0)
  int npasses;             // light-sample passes per diffuse hit
  size_t n;
  int nranks, me;
  int rows_per_block;      // home rank of image row y = (y / rows_per_block) % nranks
  size_t o_keys, o_cap, o_occ;   // offsets of THIS wave's key array / attribute array / occlusion bytes in a region
  PeerPtrs peers;
  unsigned int* ctl;       // CTL_* words
  // the wave's samples: NULL = all n samples of the tile (eye rays), else the compact list of the samples that
  // bounced (written by the previous wave's shading kernel; same set, rank-specific order)
  const uint32_t* src;
  const unsigned int* src_n;
};

__device__ __forceinline__ uint32_t p2p_pixid(const P2PWave& P, size_t i, int* row) {
  const uint32_t pix = (uint32_t)(i / (size_t)P.spp);
  const uint32_t x0 = pix % (uint32_t)P.tile.w, y0 = pix / (uint32_t)P.tile.w;
  *row = P.tile.y + (int)y0;
  return (uint32_t)(*row * P.image_w + P.tile.x + (int)x0);
}
__device__ __forceinline__ int p2p_home(const P2PWave& P, int row) { return (row / P.rows_per_block) % P.nranks; }
__device__ __forceinline__ uint32_t p2p_src_count(const P2PWave& P) { return P.src ? *P.src_n : (uint32_t)P.n; }
__global__ void __launch_bounds__(256) p2p_fill_kernel(float* __restrict__ w, size_t n, float v) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) w[i] = v;
}
__global__ void p2p_reset_kernel(unsigned int* ctl, int kb) {
  ctl[CTL_WORK1] = 0u; ctl[CTL_WORK3] = 0u; ctl[CTL_RET] = 0u; ctl[CTL_LIVE0 + (kb ^ 1)] = 0u;
}

// K1: closest hit over MY resident domains; a hit is merged into every rank's key array with a 64-bit atomicMin
// (VBuf::compositeTbuf's MPI_Allreduce(MIN), insitu_vbuf.h:109-117, done by the tracing kernel itself over NVLink).
// Persistent CTAs (the domain table is staged once), 32 samples claimed at a time; rays that enter none of MY domains
// (top-level any-test) never occupy a lane: the others are compacted through a per-warp shared-memory queue, the same
// scheme as the fused kernel (trace.cu).
__global__ void __launch_bounds__(P2P_THREADS, P2P_MINBLOCKS) p2p_trace_kernel(P2PWave P, const spraygpu_ray* __restrict__ rays) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(P.table, P.nd, P.tlas, P.tlas_n, smem);
  uint32_t* q = (uint32_t*)((char*)smem + ((dom_smem_bytes(P.nd, P.tlas_n, 0) + 15) & ~(size_t)15)) + (size_t)(threadIdx.x >> 5) * 7 * P2P_QCAP;
  const int lane = threadIdx.x & 31;
  const uint32_t nsrc = p2p_src_count(P);
  int count = 0;
  bool more = true;
  for (;;) {
    if (count < 32 && more) {
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(P.ctl + CTL_WORK1, 32u);
      base = __shfl_sync(0xffffffffu, base, 0);
      if (base >= nsrc) { more = false; continue; }
      const uint32_t k = base + lane;
      bool go = false;
      uint32_t i = 0;
      vec3 o = mk3(0, 0, 0), d = mk3(0, 0, 1);
      if (k < nsrc) {
        i = P.src ? P.src[k] : k;
        const spraygpu_ray ry = load_ray(rays + i);
        o = mk3(ry.ox, ry.oy, ry.oz); d = mk3(ry.dx, ry.dy, ry.dz);
        go = tlas_any(D, make_ray_ctx(o, d, SPRAY_RAY_EPS));
      }
      const uint32_t b = __ballot_sync(0xffffffffu, go);
      if (go) {
        uint32_t* e = q + count + __popc(b & ((1u << lane) - 1u));
        e[0 * P2P_QCAP] = __float_as_uint(o.x); e[1 * P2P_QCAP] = __float_as_uint(o.y); e[2 * P2P_QCAP] = __float_as_uint(o.z);
        e[3 * P2P_QCAP] = __float_as_uint(d.x); e[4 * P2P_QCAP] = __float_as_uint(d.y); e[5 * P2P_QCAP] = __float_as_uint(d.z);
        e[6 * P2P_QCAP] = i;
      }
      count += __popc(b);
      __syncwarp();
      continue;
    }
    if (count == 0) break;
    const int n = count < 32 ? count : 32;
    count -= n;
    const bool act = lane < n;
    const uint32_t* e = q + count + (act ? lane : 0);
    const vec3 o = mk3(__uint_as_float(e[0 * P2P_QCAP]), __uint_as_float(e[1 * P2P_QCAP]), __uint_as_float(e[2 * P2P_QCAP]));
    const vec3 d = mk3(__uint_as_float(e[3 * P2P_QCAP]), __uint_as_float(e[4 * P2P_QCAP]), __uint_as_float(e[5 * P2P_QCAP]));
    const uint32_t i = e[6 * P2P_QCAP];
    __syncwarp();
    SceneHit h;
    const bool hit = scene_trace_warp<false>(P.table, D, act, o, d, false, h, nullptr);
    if (act && hit) {
      const unsigned long long key = make_key(h.t, h.dom, h.prim);
      for (int r = 0; r < P.nranks; ++r)
        atomicMin_system((unsigned long long*)(P.peers.base[r] + P.o_keys) + i, key);
    }
  }
  grid_signal_peers(P.ctl + CTL_DONE, P.peers, P.nranks, P.me);
}

// K2: TriMeshBuffer::updateIntersection by the rank that holds the winning domain; 16 B (Ns.xyz, colour) to every rank.
// (u, v) are re-derived from (ray, prim) with the S1 arithmetic on the same vertex values: bit-identical to the traversal's.
__global__ void __launch_bounds__(128) p2p_capture_kernel(P2PWave P, const spraygpu_ray* __restrict__ rays) {
  const uint32_t nsrc = p2p_src_count(P);
  const unsigned long long* keys = (const unsigned long long*)(P.peers.base[P.me] + P.o_keys);
  for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < nsrc; k += (size_t)gridDim.x * blockDim.x) {
    const uint32_t i = P.src ? P.src[k] : (uint32_t)k;
    const unsigned long long key = keys[i];
    if (key == SPRAY_KEY_MISS) continue;
    const DomainEntry* e = P.table + key_domain(key);
    if (!e->resident) continue;
    const spraygpu_ray ry = load_ray(rays + i);
    const uint32_t prim = key_prim(key);
    const uint32_t i0 = e->faces[3 * (size_t)prim], i1 = e->faces[3 * (size_t)prim + 1], i2 = e->faces[3 * (size_t)prim + 2];
    const vec3 v0 = mk3(e->verts[3 * (size_t)i0], e->verts[3 * (size_t)i0 + 1], e->verts[3 * (size_t)i0 + 2]);
    const vec3 v1 = mk3(e->verts[3 * (size_t)i1], e->verts[3 * (size_t)i1 + 1], e->verts[3 * (size_t)i1 + 2]);
    const vec3 v2 = mk3(e->verts[3 * (size_t)i2], e->verts[3 * (size_t)i2 + 1], e->verts[3 * (size_t)i2 + 2]);
    TriHit th; th.t = th.u = th.v = 0.f;
    tri_test(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), v0, gsub(v0, v1), gsub(v2, v0), ry.tmin, &th);
    uint32_t col; vec3 Ns;
    fetch_attributes(e, prim, th.u, th.v, &col, &Ns);
    const uint4 rec = make_uint4(__float_as_uint(Ns.x), __float_as_uint(Ns.y), __float_as_uint(Ns.z), col);
    for (int r = 0; r < P.nranks; ++r) ((uint4*)(P.peers.base[r] + P.o_cap))[i] = rec;
  }
  grid_signal_peers(P.ctl + CTL_DONE, P.peers, P.nranks, P.me);
}

// K3: EVERY rank shades every hit from (reduced key, exchanged attributes) — ShaderPt / ShaderAo
// (insitu_shader_pt.h:83-230, insitu_shader_ao.h) — so the shadow rays, the bounce ray and the radiance are identical on
// all ranks without travelling.  Each rank tests the shadow rays against ITS domains and reports an occlusion to the
// sample's home rank (VBuf::compositeObuf's MPI_Allreduce(BOR), insitu_vbuf.h:119-129, as single-byte peer stores).
// Persistent CTAs; the hit samples are compacted through a per-warp index queue, so shading and shadow tracing run with
// full warps although only a fraction of the wave's samples hit anything.  Samples that bounce are appended to the next
// wave's list; the home rank appends (sample, pass, radiance) of its own shadow rays to the retire list.
__global__ void __launch_bounds__(P2P_THREADS, P2P_MINBLOCKS)
    p2p_shade_kernel(P2PWave P, const spraygpu_ray* __restrict__ rays, const float* __restrict__ w, spraygpu_ray* __restrict__ rays_next,
                     float* __restrict__ w_next, uint32_t* __restrict__ next_list, unsigned int* __restrict__ next_n,
                     uint4* __restrict__ ret, unsigned long long* __restrict__ counters) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(P.table, P.nd, P.tlas, P.tlas_n, smem);
  uint32_t* q = (uint32_t*)((char*)smem + ((dom_smem_bytes(P.nd, P.tlas_n, 0) + 15) & ~(size_t)15)) + (size_t)(threadIdx.x >> 5) * P2P_QCAP;
  const int lane = threadIdx.x & 31;
  const uint32_t nsrc = p2p_src_count(P);
  const unsigned long long* keys = (const unsigned long long*)(P.peers.base[P.me] + P.o_keys);
  const uint4* cap = (const uint4*)(P.peers.base[P.me] + P.o_cap);
  const bool pt = P.shade.shader == SPRAYGPU_SHADER_PT;
  const int nl = pt ? P.shade.nlights : 1;
  const int next_depth = P.depth + 1;
  uint32_t n_hit = 0, n_shadow = 0, n_rad = 0;
  int count = 0;
  bool more_src = true;
  for (;;) {
    if (count < 32 && more_src) {
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(P.ctl + CTL_WORK3, 32u);
      base = __shfl_sync(0xffffffffu, base, 0);
      if (base >= nsrc) { more_src = false; continue; }
      const uint32_t k = base + lane;
      bool go = false;
      uint32_t i = 0;
      if (k < nsrc) {
        i = P.src ? P.src[k] : k;
        go = keys[i] != SPRAY_KEY_MISS;
      }
      const uint32_t b = __ballot_sync(0xffffffffu, go);
      if (go) q[count + __popc(b & ((1u << lane) - 1u))] = i;
      count += __popc(b);
      __syncwarp();
      continue;
    }
    if (count == 0) break;
    const int n = count < 32 ? count : 32;
    count -= n;
    const bool hit = lane < n;
    const uint32_t i = q[count + (hit ? lane : 0)];
    __syncwarp();
    const unsigned long long key = keys[i];
    const spraygpu_ray ry = load_ray(rays + i);
    const vec3 ro = mk3(ry.ox, ry.oy, ry.oz), rd = mk3(ry.dx, ry.dy, ry.dz);
    int row = 0;
    const uint32_t pixid = p2p_pixid(P, i, &row);
    const int home = p2p_home(P, row);
    const bool mine = hit && home == P.me;
    const uint32_t samid = i;
    HitCtx c;
    c.pos = ro; c.kd = mk3(0, 0, 0); c.nff = mk3(0, 0, 1); c.wo = gneg(rd);
    c.cos_theta_i = 0.f; c.entering = false; c.mtl = SPRAYGPU_MTL_DIFFUSE; c.eta0 = c.eta1 = 1.f;
    vec3 Lin = mk3(0, 0, 0), bd = mk3(0, 0, 0), bw = mk3(0, 0, 0);
    bool bounce = false;
    if (hit) {
      const uint4 rec = cap[i];
      const DomainEntry* e = P.table + key_domain(key);        // material lives in the table, resident or not
      const vec3 normal = mk3(__uint_as_float(rec.x), __uint_as_float(rec.y), __uint_as_float(rec.z));
      c.pos = hit_position(ro, rd, key_t(key));
      c.kd = unpack_color(rec.w);
      c.cos_theta_i = gdot(c.wo, normal);
      c.entering = c.cos_theta_i > 0.0f;
      c.nff = gnormalize(c.entering ? normal : gneg(normal));
      c.mtl = e->mtl_type; c.eta0 = e->mtl[0]; c.eta1 = e->mtl[1];
      Lin = mk3(w[3 * (size_t)i], w[3 * (size_t)i + 1], w[3 * (size_t)i + 2]);
      if (mine) ++n_hit;
      if (next_depth < P.shade.bounces) {
        bool em1, em2; vec3 d1 = mk3(0, 0, 1), w1 = mk3(0, 0, 0), d2 = mk3(0, 0, 1), w2 = mk3(0, 0, 0);
        shade_secondary<true>(c, Lin, samid, next_depth, &em1, &d1, &w1, &em2, &d2, &w2);
        // glass domains are rejected at launch, so at most one of the two is live
        if (em1) { bd = d1; bw = w1; bounce = true; }
        else if (em2) { bd = d2; bw = w2; bounce = true; }
        if (bounce && mine) ++n_rad;
      }
    }
    // next wave (identical set on every rank): origin = hit position, direction = bounce
    {
      const uint32_t slot = warp_append(next_n, bounce);
      if (bounce) {
        next_list[slot] = i;
        spraygpu_ray nx;
        nx.ox = c.pos.x; nx.oy = c.pos.y; nx.oz = c.pos.z; nx.tmin = SPRAY_RAY_EPS;
        nx.dx = bd.x; nx.dy = bd.y; nx.dz = bd.z; nx.tmax = INFINITY;
        store_ray(rays_next + i, nx);
        w_next[3 * (size_t)i] = bw.x; w_next[3 * (size_t)i + 1] = bw.y; w_next[3 * (size_t)i + 2] = bw.z;
      }
    }
    uint32_t ls = rng_init1(samid * (uint32_t)next_depth);     // light_sampler (insitu_shader_pt.h:115)
    int l = 0, k = 0;
    bool more = hit && c.mtl == SPRAYGPU_MTL_DIFFUSE && nl > 0;
    for (int p = 0; p < P.npasses; ++p) {
      if (!__any_sync(0xffffffffu, more)) break;
      bool emit = false; vec3 wi = mk3(0, 0, 1), Lr = mk3(0, 0, 0);
      if (more) emit = shade_light_pass<true>(P.shade, pt, nl, c, Lin, pixid, l, k, ls, &wi, &Lr, &more);
      SceneHit hs;
      const bool oc = scene_trace_warp<false>(P.table, D, emit, c.pos, wi, true, hs, nullptr);
      const size_t j = (size_t)p * P.n + i;
      if (emit && oc) ((uint8_t*)(P.peers.base[home] + P.o_occ))[j] = 1;
      const bool keep = mine && emit;
      if (keep) ++n_shadow;
      const uint32_t slot = warp_append(P.ctl + CTL_RET, keep);
      if (keep) ret[slot] = make_uint4((uint32_t)j, __float_as_uint(Lr.x), __float_as_uint(Lr.y), __float_as_uint(Lr.z));
    }
  }
  unsigned long long v[3] = {n_rad, n_shadow, n_hit};
  const int slots[3] = {2, 3, 7};
#pragma unroll
  for (int qq = 0; qq < 3; ++qq) {
    unsigned long long x = v[qq];
#pragma unroll
    for (int off = 16; off; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
    if (lane == 0 && x) atomicAdd(counters + slots[qq], x);
  }
  grid_signal_peers(P.ctl + CTL_DONE, P.peers, P.nranks, P.me);
}

// K4: TContext::retireShadows -> HdrImage::add on the sample's home rank (entries: this rank's emitted shadow rays)
__global__ void __launch_bounds__(256) p2p_retire_kernel(P2PWave P, const uint4* __restrict__ ret, float* __restrict__ fb) {
  const unsigned int nret = P.ctl[CTL_RET];
  const uint8_t* occ = (const uint8_t*)(P.peers.base[P.me] + P.o_occ);
  const float scale = (float)(1.0 / (double)P.spp);
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < nret; e += (size_t)gridDim.x * blockDim.x) {
    const uint4 r = ret[e];
    if (occ[r.x]) continue;
    int row;
    const uint32_t pixid = p2p_pixid(P, (size_t)r.x % P.n, &row);
    fb_add(fb, pixid, scale, mk3(__uint_as_float(r.y), __uint_as_float(r.z), __uint_as_float(r.w)));
  }
}

}  // namespace spraygpu

using namespace spraygpu;

static void p2p_free_local(P2PState* S) {
  for (int k = 0; k < 2; ++k) {
    if (S->rays[k]) cudaFree(S->rays[k]);
    if (S->w[k]) cudaFree(S->w[k]);
    S->rays[k] = nullptr; S->w[k] = nullptr;
  }
  for (int k = 0; k < 2; ++k) { if (S->list[k]) cudaFree(S->list[k]); S->list[k] = nullptr; }
  if (S->ret) cudaFree(S->ret);
  S->ret = nullptr;
}
// registered with the communicator: runs in spraygpu_comm_destroy (collective)
static void p2p_destroy(Comm* c, void* p) {
  P2PState* S = (P2PState*)p;
  if (!S) return;
  cudaSetDevice(comm_device(c)->cuda_device);
  cudaStreamSynchronize(comm_device(c)->stream);
  if (S->local) { comm_host_barrier(c); comm_shared_free(c, S->local, &S->mapped); }
  p2p_free_local(S);
  if (S->d_ctl) cudaFree(S->d_ctl);
  if (S->h_err) cudaFreeHost(S->h_err);
  delete S;
}

// collective: every rank calls with the same (n, npasses)
static int p2p_ensure(Comm* c, P2PState* S, size_t n, int npasses) {
  Device* dev = comm_device(c);
  if (!S->d_ctl) {
    SG_CUDA(cudaMalloc((void**)&S->d_ctl, sizeof(unsigned int) * CTL_WORDS));
    SG_CUDA(cudaMemset(S->d_ctl, 0, sizeof(unsigned int) * CTL_WORDS));
    SG_CUDA(cudaHostAlloc((void**)&S->h_err, sizeof(unsigned int), cudaHostAllocMapped));
    *S->h_err = 0u;
    SG_CUDA(cudaHostGetDevicePointer((void**)&S->h_err_dev, S->h_err, 0));
  }
  if (n <= S->cap_n && npasses <= S->cap_passes) return SPRAYGPU_OK;
  SG_CUDA(cudaStreamSynchronize(dev->stream));
  if (S->local) {
    int rc = comm_host_barrier(c);           // nobody is still storing into a region that is about to be replaced
    if (rc) return rc;
    comm_shared_free(c, S->local, &S->mapped);
    S->local = nullptr;
  }
  p2p_free_local(S);
  const size_t cn = std::max(n, S->cap_n);
  const int cp = std::max(npasses, std::max(S->cap_passes, 1));
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  size_t off = P2P_FLAG_BYTES;
  S->o_keys[0] = off; off += al(cn * 8);
  S->o_keys[1] = off; off += al(cn * 8);
  S->o_cap = off; off += al(cn * 16);
  S->o_occ = off; off += al(cn * (size_t)cp);
  S->region_bytes = off;
  int rc = comm_shared_alloc(c, off, &S->local, &S->mapped, -1);   // zero-filled: flag = 0
  if (rc) return rc;
  S->seq = 0;
  for (int k = 0; k < 2; ++k) {
    SG_CUDA(cudaMalloc((void**)&S->rays[k], cn * sizeof(spraygpu_ray)));
    SG_CUDA(cudaMalloc((void**)&S->w[k], cn * 12));
  }
  for (int k = 0; k < 2; ++k) SG_CUDA(cudaMalloc((void**)&S->list[k], cn * sizeof(uint32_t)));
  SG_CUDA(cudaMalloc((void**)&S->ret, cn * sizeof(uint4) * (size_t)cp));
  S->cap_n = cn; S->cap_passes = cp;
  return comm_host_barrier(c);               // every region exists and is mapped before the first signal
}

static inline int p2p_grid(size_t n, int t) { return (int)((n + t - 1) / t); }

extern "C" int spraygpu_trace_tile_insitu_p2p(spraygpu_device dev_h, spraygpu_comm comm_h, const spraygpu_camera* cam,
                                              const spraygpu_shade* shade, spraygpu_tile tile, float* fb, int rows_per_block,
                                              spraygpu_stats* stats) {
  SG_CHECK(dev_h && comm_h && cam && shade && fb, SPRAYGPU_ERR_ARG, "trace_tile_insitu_p2p: NULL argument");
  Comm* comm = (Comm*)comm_h;
  Device* dev = (Device*)dev_h;
  SG_CHECK(comm_device(comm) == dev, SPRAYGPU_ERR_ARG, "trace_tile_insitu_p2p: communicator belongs to another device");
  SG_CHECK(comm_size(comm) <= P2P_MAX_RANKS, SPRAYGPU_ERR_LIMIT, "trace_tile_insitu_p2p: %d ranks (limit %d)", comm_size(comm), P2P_MAX_RANKS);
  SG_CHECK(rows_per_block >= 1, SPRAYGPU_ERR_ARG, "trace_tile_insitu_p2p: rows_per_block must be >= 1");
  SG_CHECK(shade->pixel_samples >= 1 && shade->bounces >= 1, SPRAYGPU_ERR_ARG, "trace_tile_insitu_p2p: pixel_samples and bounces must be >= 1");
  SG_CHECK(shade->nlights >= 0 && shade->nlights <= SPRAYGPU_MAX_LIGHTS, SPRAYGPU_ERR_LIMIT, "trace_tile_insitu_p2p: %d lights", shade->nlights);
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_AO || shade->nlights > 0, SPRAYGPU_ERR_ARG, "trace_tile_insitu_p2p: the PT shader needs a light");
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_PT || shade->ao_samples > 0, SPRAYGPU_ERR_ARG, "trace_tile_insitu_p2p: ao_samples must be > 0");
  SG_CHECK(tile.x >= 0 && tile.y >= 0 && tile.w >= 0 && tile.h >= 0 && tile.x + tile.w <= cam->image_w &&
               tile.y + tile.h <= cam->image_h, SPRAYGPU_ERR_ARG, "trace_tile_insitu_p2p: tile outside the image");
  SG_CUDA(cudaSetDevice(dev->cuda_device));
  {
    cudaPointerAttributes a;
    const bool devptr = cudaPointerGetAttributes(&a, fb) == cudaSuccess && (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    SG_CHECK(devptr, SPRAYGPU_ERR_ARG, "trace_tile_insitu_p2p: fb must be a device pointer");
  }
  for (int i = 0; i < dev->ndomains; ++i)
    SG_CHECK(dev->h_domains[i].mtl_type != SPRAYGPU_MTL_GLASS, SPRAYGPU_ERR_LIMIT,
             "trace_tile_insitu_p2p: glass domains (two secondary rays per hit) are not supported on the in-situ path; "
             "the reference CHECK-fails on them too (insitu_shader_pt.h)");
  void (**free_slot)(Comm*, void*) = nullptr;
  void** slot = comm_p2p_slot(comm, &free_slot, 0);
  if (!*slot) { *slot = new P2PState(); *free_slot = p2p_destroy; }
  P2PState* S = (P2PState*)*slot;
  { int erc = p2p_take_error(S->h_err, "trace_tile_insitu_p2p"); if (erc) return erc; }
  int rc = upload_domain_table(dev);
  if (rc) return rc;
  const size_t n = (size_t)tile.w * tile.h * shade->pixel_samples;
  cudaStream_t st = dev->stream;
  SG_CUDA(cudaMemsetAsync(dev->d_counters, 0, sizeof(unsigned long long) * 8, st));
  if (n) {
    int npasses = 0;
    if (shade->shader == SPRAYGPU_SHADER_PT)
      for (int l = 0; l < shade->nlights; ++l) npasses += shade->lights[l].type == SPRAYGPU_LIGHT_DIFFUSE ? shade->ao_samples : 1;
    else npasses = shade->ao_samples;
    rc = p2p_ensure(comm, S, n, npasses);
    if (rc) return rc;
    const int nranks = comm_size(comm), me = comm_rank(comm);
    SG_CHECK(n * (size_t)(npasses > 0 ? npasses : 1) < ((size_t)1 << 32), SPRAYGPU_ERR_LIMIT,
             "trace_tile_insitu_p2p: %zu samples x %d light passes do not fit 32-bit list entries; use smaller blocking tiles", n, npasses);
    P2PWave P;
    P.table = dev->d_domains; P.nd = dev->ndomains; P.tlas = dev->d_tlas; P.tlas_n = dev->tlas_n;
    P.shade = *shade; P.tile = tile; P.image_w = cam->image_w; P.spp = shade->pixel_samples; P.npasses = npasses; P.n = n;
    P.nranks = nranks; P.me = me; P.rows_per_block = rows_per_block; P.o_cap = S->o_cap; P.o_occ = S->o_occ; P.ctl = S->d_ctl;
    for (int r = 0; r < P2P_MAX_RANKS; ++r) P.peers.base[r] = r < nranks ? (char*)S->mapped[(size_t)r] : nullptr;
    char* mine = (char*)S->local;
    auto barrier_wait = [&]() {
      ++S->seq;
      p2p_wait_kernel<<<1, 1, 0, st>>>((volatile unsigned long long*)mine, S->seq * (unsigned long long)(nranks - 1), S->h_err_dev, p2p_timeout_cycles());
      dev->launches++;
    };
    const size_t dsmem = (dom_smem_bytes(dev->ndomains, dev->tlas_n, 0) + 15) & ~(size_t)15;
    const size_t smem_trace = dsmem + sizeof(uint32_t) * 7 * P2P_QCAP * (P2P_THREADS / 32);
    const size_t smem_shade = dsmem + sizeof(uint32_t) * P2P_QCAP * (P2P_THREADS / 32);
    if (S->grid_smem != dsmem) {
      SG_DYN_SMEM(p2p_trace_kernel, smem_trace);
      SG_DYN_SMEM(p2p_shade_kernel, smem_shade);
      int b1 = 0, b3 = 0;
      SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b1, p2p_trace_kernel, P2P_THREADS, smem_trace));
      SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b3, p2p_shade_kernel, P2P_THREADS, smem_shade));
      SG_CHECK(b1 > 0 && b3 > 0, SPRAYGPU_ERR_LIMIT, "trace_tile_insitu_p2p: the wave kernels do not fit an SM (%zu B shared)", smem_trace);
      S->grid_trace = b1 * dev->sm_count; S->grid_shade = b3 * dev->sm_count; S->grid_smem = dsmem;
    }
    const int grid_lin = dev->sm_count * 8;
    PhaseTimer T(st);
    rc = launch_eye_rays(dev, cam, tile, shade->pixel_samples, SPRAYGPU_EYE_SEED_INSITU, S->rays[0]);   // insitu_ray.h:103-182
    if (rc) return rc;
    p2p_fill_kernel<<<p2p_grid(3 * n, 256), 256, 0, st>>>(S->w[0], 3 * n, 1.0f);
    dev->launches++;
    SG_CUDA(cudaMemsetAsync(mine + S->o_keys[0], 0xFF, n * 8, st));
    p2p_signal_kernel<<<1, 1, 0, st>>>(P.peers, nranks, me);       // everyone's first key array is reset
    dev->launches++;
    T.mark("eye+init");
    barrier_wait();
    T.mark("bar");
    int cur = 0;
    for (int depth = 0; depth < shade->bounces; ++depth) {
      const int kb = depth & 1;
      P.depth = depth; P.o_keys = S->o_keys[kb];
      P.src = depth ? S->list[kb] : nullptr;                       // wave 0: every sample of the tile
      P.src_n = S->d_ctl + CTL_LIVE0 + kb;
      SG_CUDA(cudaMemsetAsync(mine + S->o_keys[kb ^ 1], 0xFF, n * 8, st));          // the NEXT wave's keys
      if (npasses) SG_CUDA(cudaMemsetAsync(mine + S->o_occ, 0, n * (size_t)npasses, st));
      p2p_reset_kernel<<<1, 1, 0, st>>>(S->d_ctl, kb);
      p2p_trace_kernel<<<S->grid_trace, P2P_THREADS, smem_trace, st>>>(P, S->rays[cur]);      // K1
      T.mark("trace");
      barrier_wait();
      T.mark("bar");
      p2p_capture_kernel<<<grid_lin, 128, 0, st>>>(P, S->rays[cur]);                           // K2
      T.mark("capture");
      barrier_wait();
      T.mark("bar");
      p2p_shade_kernel<<<S->grid_shade, P2P_THREADS, smem_shade, st>>>(P, S->rays[cur], S->w[cur], S->rays[cur ^ 1], S->w[cur ^ 1],
                                                                       S->list[kb ^ 1], S->d_ctl + CTL_LIVE0 + (kb ^ 1), S->ret,
                                                                       dev->d_counters);      // K3
      T.mark("shade");
      barrier_wait();
      T.mark("bar");
      if (npasses) {
        p2p_retire_kernel<<<grid_lin, 256, 0, st>>>(P, S->ret, fb);                            // K4
        dev->launches++;
        T.mark("retire");
      }
      dev->launches += 4;
      cur ^= 1;
    }
    T.report("insitu_p2p", me);
    SG_CUDA(cudaGetLastError());
  }
  if (stats) {
    // every rank counted the samples it is home to; eye rays (n) are identical on every rank
    rc = comm_allreduce_u64_sum(comm, dev->d_counters, 8);
    if (rc) return rc;
    unsigned long long* h = (unsigned long long*)dev->pinned;
    SG_CUDA(cudaMemcpyAsync(h, dev->d_counters, sizeof(unsigned long long) * 8, cudaMemcpyDeviceToHost, st));
    SG_CUDA(cudaStreamSynchronize(st));
    { int erc = p2p_take_error(S->h_err, "trace_tile_insitu_p2p"); if (erc) return erc; }
    stats->eye_rays = n; stats->radiance_rays = n + h[2]; stats->shadow_rays = h[3]; stats->domain_tests = 0;
    stats->nodes_visited = 0; stats->tris_tested = 0; stats->hits = h[7];
  }
  return SPRAYGPU_OK;
}
