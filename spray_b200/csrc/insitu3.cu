// insitu3.cu — the in-situ (domain-partitioned) blocking tile with TRUE RAY FORWARDING: spraygpu_trace_tile_insitu_fwd.
//
// Reference: insitu::MultiThreadTracer (src/insitu/insitu_multithread_tracer.inl:135-195, :313-442) — a rank traces the
// rays of ITS image share in its own domains and SENDS every ray that also enters a remote domain to that domain's
// owner (insitu::Comm, src/insitu/insitu_comm.inl:51-92: per-destination MsgHeader + Ray[] messages over MPI_Isend /
// MPI_Iprobe), owners trace what they receive and the t-buffer / occlusion buffer decide which speculative result
// survives (insitu_vbuf.h:74-152); WorkStats counts outstanding work for termination (insitu_work_stats.cc:34-85).
//
// Here the messages are 32-byte ray records appended straight into the OWNER's inbox over NVLink /
// NVSwitch peer memory (CUDA IPC): one warp-aggregated system-scope atomicAdd on the owner's counter + two 16-byte
// stores per ray, from inside the kernel that generated the ray.  Only rays that ENTER a domain of that owner travel
// (exact slab test, aabb.h:139-169, against every domain box; 0.67 domain tests per ray on C2), and every sample is
// shaded ONCE, on its home rank (interleaved blocks of image rows) — the dense variant (insitu2.cu) regenerates,
// culls and shades every sample on every rank.
//   F1 route   : home rank: eye rays of its samples -> inbox of every rank that owns an entered domain (itself included)
//   F2 serve   : owner: trace the inbox against ITS domains.  radiance ray: 64-bit atomicMin of the packed
//                (t | domain | prim) key into the HOME rank's key array, remember the record as a candidate winner;
//                shadow ray: store 1 into the home rank's occlusion byte when occluded
//   F3 capture : owner: for each candidate whose key is the final one, TriMeshBuffer::updateIntersection
//                (colour + shading normal, 16 B) into the home rank's attribute array
//   F4 shade   : home: ShaderPt / ShaderAo on (key, attributes); the shadow rays and the bounce ray are routed into the
//                owners' inboxes of the NEXT wave (the reference's speculation: shadow rays of depth d travel with the
//                radiance rays of depth d + 1), retire list kept at home
//   F5 retire  : home: HdrImage::add of the unoccluded shadow rays of the previous wave (occlusion final after F2)
// Three device-side barriers per wave (after F2, F3, F4), none of them a host round trip; termination = the fixed
// number of waves (`bounces`), an empty wave costs three near-empty launches.
// Results are those of the fused kernel / the dense variant: same keys (lexicographic min), same shading arithmetic.
#include <algorithm>
#include <vector>

#include "internal.cuh"
#include "traverse.cuh"
#include "shade.cuh"
#include "peer.cuh"

namespace spraygpu {

struct FwdState {
  size_t cap_n = 0;
  int cap_passes = 0;
  size_t seg_cap = 0;          // records per (source rank) segment of an inbox
  void* local = nullptr;       // the peer-mapped region of THIS rank
  std::vector<void*> mapped;
  size_t o_inbox[2] = {0, 0}, o_keys[2] = {0, 0}, o_cap = 0, o_occ = 0;
  unsigned long long seq = 0;
  // rank-local
  spraygpu_ray* rays[2] = {nullptr, nullptr};
  float* w[2] = {nullptr, nullptr};
  uint32_t* list[2] = {nullptr, nullptr};
  uint4* ret = nullptr;                    // (pass * n + sample, radiance)
  uint32_t* cand_idx = nullptr;            // inbox record index of a candidate winner
  unsigned long long* cand_key = nullptr;  // the key it merged
  unsigned int* d_ctl = nullptr;
  float* d_route = nullptr;                // rank boxes | rank ranges | domain permutation (see RouteTab)
  int grid_serve = 0, grid_shade = 0;
  size_t grid_smem = (size_t)-1;
  unsigned int* h_err = nullptr;
  unsigned int* h_err_dev = nullptr;
};

// control words (rank-local device memory)
// FC_PUSH0 / FC_PUSH1: records THIS rank has appended to inbox 0 / 1 of every destination (16 words each).  A
// destination's inbox is cut into one segment per SOURCE rank, so slots are handed out by a LOCAL atomic — the first
// version took them from a counter in the destination's memory: one NVLink round trip (~2-3 us) per (warp,
// destination), and route0 / shade were bound by exactly that latency (130 / 123 us for 0.5 M rays at 8 GPUs).
enum { FC_DONE = 0, FC_WORK_ROUTE, FC_WORK_SERVE, FC_WORK_SHADE, FC_RET, FC_CAND, FC_LIVE0, FC_LIVE1, FC_OVERFLOW, FC_PUSH0 = 16,
       FC_PUSH1 = 32, FC_WORDS = 48 };
// words at P2P_FLAG_BYTES/2 of a region: records per source segment of inbox 0 (16 words) and inbox 1 (16 words), written
// by the sources when they have finished appending (fwd_publish_kernel)
#define FWD_CTR_OFFSET 128
#define FWD_THREADS 128
#define FWD_MINBLOCKS 4
#define FWD_QCAP 64

// routing table in device memory / shared memory: per rank the box around its domains, then [begin, end) into `perm`,
// the domain ids sorted by owner
struct RouteTab {
  const float* rank_box;   // nranks x 6
  const int* begin;        // nranks + 1
  const int* perm;         // nd
};
__host__ __device__ inline size_t route_words(int nranks, int nd) { return (size_t)nranks * 6 + (size_t)nranks + 1 + (size_t)nd; }

struct FwdWave {
  const DomainEntry* table;
  int nd;
  const TlasNode* tlas;
  int tlas_n;
  const float* route;      // device copy of the routing table
  spraygpu_shade shade;
  spraygpu_tile tile;
  CameraDev cam;
  int image_w, spp;
  int depth, npasses;
  size_t n;
  int nranks, me, rows_per_block;
  size_t o_inbox_cur, o_inbox_next, o_keys, o_keys_next, o_cap, o_occ;
  int inbox_cur, inbox_next;   // which counter word
  uint32_t seg_cap;
  PeerPtrs peers;
  unsigned int* ctl;
  const uint32_t* src;         // wave > 0: the live home samples
  const unsigned int* src_n;
};

__device__ __forceinline__ uint32_t fwd_pixid(const FwdWave& P, size_t i, int* row) {
  const uint32_t pix = (uint32_t)(i / (size_t)P.spp);
  const uint32_t x0 = pix % (uint32_t)P.tile.w, y0 = pix / (uint32_t)P.tile.w;
  *row = P.tile.y + (int)y0;
  return (uint32_t)(*row * P.image_w + P.tile.x + (int)x0);
}
__device__ __forceinline__ int fwd_home(const FwdWave& P, int row) { return (row / P.rows_per_block) % P.nranks; }

__device__ __forceinline__ RouteTab load_route_smem(const float* __restrict__ route, int nranks, int nd, float* smem) {
  const int nw = (int)route_words(nranks, nd);
  for (int i = threadIdx.x; i < nw; i += blockDim.x) smem[i] = route[i];
  __syncthreads();
  RouteTab R;
  R.rank_box = smem; R.begin = (const int*)(smem + 6 * nranks); R.perm = R.begin + nranks + 1;
  return R;
}

// ranks that own a domain this ray enters: the reference's exact slab test (aabb.h:139-169) with tnear = 1e-3, the same
// test the owner's traversal applies before it opens the domain (TlasSource), so nothing is sent in vain and nothing
// missed.  *primary = the rank that owns the NEAREST entered domain (smallest entry distance; ties -> lowest rank): it
// traces first, the others trace afterwards against the hit it found (insitu::Isector::isectWithoutFarDomain,
// src/insitu/insitu_isector.h:165-224: "all but ... nearer than the known hit t").
__device__ __forceinline__ uint32_t route_mask(const float* __restrict__ boxes, const RouteTab& R, int nranks, vec3 o, vec3 d, int* primary) {
  const RayCtx r = make_ray_ctx(o, d, SPRAY_RAY_EPS);
  const float ix = div_rn(1.0f, d.x), iy = div_rn(1.0f, d.y), iz = div_rn(1.0f, d.z);
  uint32_t mask = 0;
  float best = INFINITY;
  int prim = -1;
  for (int rk = 0; rk < nranks; ++rk) {
    const int b = R.begin[rk], e = R.begin[rk + 1];
    if (b == e) continue;
    const float* rb = R.rank_box + 6 * rk;
    float t;
    if (!tlas_box_hit(rb[0], rb[1], rb[2], rb[3], rb[4], rb[5], r, &t)) continue;   // conservative
    float near = INFINITY;
    bool any = false;
    for (int k = b; k < e; ++k) {
      float tmin;
      if (domain_slab_inv(boxes + 6 * R.perm[k], o, ix, iy, iz, SPRAY_RAY_EPS, INFINITY, &tmin)) { any = true; near = fminf(near, tmin); }
    }
    if (any) {
      mask |= 1u << rk;
      if (prim < 0 || near < best) { best = near; prim = rk; }
    }
  }
  *primary = prim;
  return mask;
}

// ALL 32 lanes call.  Appends this lane's ray to the inbox of every rank in `mask` (insitu::Comm::send): one system-scope
// atomicAdd per (warp, destination), two 16-byte peer stores per record.
//   record = { ox, oy, oz, sample | dx, dy, dz, meta },
//   meta = home rank | kind << 8 (0 radiance, 1 shadow) | secondary << 9 (not the owner of the nearest domain) | pass << 16
#define FWD_META_SHADOW (1u << 8)
#define FWD_META_SECONDARY (1u << 9)
__device__ __forceinline__ void fwd_push(const FwdWave& P, size_t o_inbox, int which, uint32_t mask, int primary, vec3 o, vec3 d,
                                         uint32_t sample, uint32_t meta) {
  const int lane = threadIdx.x & 31;
  uint32_t all = __reduce_or_sync(0xffffffffu, mask);
  while (all) {
    const int rk = __ffs((int)all) - 1;
    all &= all - 1;
    const bool go = (mask >> rk) & 1u;
    const uint32_t b = __ballot_sync(0xffffffffu, go);
    const int leader = __ffs((int)b) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(P.ctl + (which ? FC_PUSH1 : FC_PUSH0) + rk, (unsigned int)__popc(b));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (go) {
      const uint32_t slot = base + __popc(b & ((1u << lane) - 1u));
      if (slot < P.seg_cap) {
        uint4* rec = (uint4*)(P.peers.base[rk] + o_inbox) + 2 * ((size_t)P.me * P.seg_cap + slot);
        rec[0] = make_uint4(__float_as_uint(o.x), __float_as_uint(o.y), __float_as_uint(o.z), sample);
        rec[1] = make_uint4(__float_as_uint(d.x), __float_as_uint(d.y), __float_as_uint(d.z), meta | (rk == primary ? 0u : FWD_META_SECONDARY));
      } else {
        P.ctl[FC_OVERFLOW] = 1u;
      }
    }
  }
}

__global__ void fwd_reset_kernel(unsigned int* ctl, int zero_push /*bit 0 / bit 1*/, int zero_live /*-1: none*/) {
  ctl[FC_WORK_ROUTE] = 0u; ctl[FC_WORK_SERVE] = 0u; ctl[FC_WORK_SHADE] = 0u; ctl[FC_CAND] = 0u;
  for (int r = 0; r < P2P_MAX_RANKS; ++r) {
    if (zero_push & 1) ctl[FC_PUSH0 + r] = 0u;
    if (zero_push & 2) ctl[FC_PUSH1 + r] = 0u;
  }
  if (zero_live >= 0) ctl[FC_LIVE0 + zero_live] = 0u;
  if (zero_push == 3) ctl[FC_OVERFLOW] = 0u;     // tile start
}
// one thread, at the end of a tile: an overflow goes into the mapped error word, so that the NEXT call reports it even
// when nobody asks for statistics (the timed frame path passes stats = NULL), and does not outlive the tile
__global__ void fwd_report_kernel(unsigned int* ctl, unsigned int* err) {
  if (ctl[FC_OVERFLOW]) { *err = *(volatile unsigned int*)err | P2P_ERR_OVERFLOW; ctl[FC_OVERFLOW] = 0u; }
}
// one thread, after a kernel that appended records: tell every destination how many records my segment of its inbox
// `which` holds, then signal the barrier (the appending kernel has completed: its peer stores are performed)
__global__ void fwd_publish_kernel(PeerPtrs P, int nranks, int me, const unsigned int* ctl, int which, uint32_t seg_cap) {
  for (int r = 0; r < nranks; ++r) {
    const unsigned int c = ctl[(which ? FC_PUSH1 : FC_PUSH0) + r];
    ((unsigned int*)(P.base[r] + FWD_CTR_OFFSET))[which * P2P_MAX_RANKS + me] = c < seg_cap ? c : seg_cap;
  }
  __threadfence_system();
  for (int r = 0; r < nranks; ++r)
    if (r != me) atomicAdd_system((unsigned long long*)P.base[r], 1ull);
}
// record index k of the concatenated segments -> physical record index (segment * seg_cap + offset)
__device__ __forceinline__ uint32_t fwd_phys(const uint32_t* seg_end, int nranks, uint32_t seg_cap, uint32_t k) {
  int sgm = 0;
  while (sgm < nranks - 1 && k >= seg_end[sgm]) ++sgm;
  return (uint32_t)sgm * seg_cap + (k - (sgm ? seg_end[sgm - 1] : 0u));
}
__global__ void fwd_zero_ret_kernel(unsigned int* ctl) { ctl[FC_RET] = 0u; }

// F1: eye rays of MY home samples (Camera::generateRay, insitu::gen{Single,Multi}SampleEyeRays, insitu_ray.h:103-182),
// kept in rays[0] for the shading pass and routed to the owners of the domains they enter.  One warp = 32 consecutive
// samples (one image row unless the row is shorter), so the home test is warp-uniform.
__global__ void __launch_bounds__(FWD_THREADS) fwd_route0_kernel(FwdWave P, spraygpu_ray* __restrict__ rays) {
  extern __shared__ float smem[];
  float* boxes = smem;
  for (int i = threadIdx.x; i < 6 * P.nd; i += blockDim.x) boxes[i] = P.table[i / 6].box[i % 6];
  const RouteTab R = load_route_smem(P.route, P.nranks, P.nd, smem + 6 * P.nd);   // syncs
  const int lane = threadIdx.x & 31;
  const size_t nchunks = (P.n + 31) / 32;
  const size_t warps = (size_t)gridDim.x * (blockDim.x >> 5);
  for (size_t c = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < nchunks; c += warps) {
    const size_t i = c * 32 + lane;
    bool go = false;
    vec3 o = mk3(P.cam.pos[0], P.cam.pos[1], P.cam.pos[2]), d = mk3(0, 0, 1);
    if (i < P.n) {
      int row;
      fwd_pixid(P, i, &row);
      go = fwd_home(P, row) == P.me;
    }
    if (!__any_sync(0xffffffffu, go)) continue;
    uint32_t mask = 0;
    int primary = -1;
    if (go) {
      const uint32_t sidx = (uint32_t)(i % (size_t)P.spp), pix = (uint32_t)(i / (size_t)P.spp);
      uint32_t samid, pixid;
      eye_sample(P.cam, P.tile, P.tile.w, 0, 0, P.spp, SPRAYGPU_EYE_SEED_INSITU, pix % (uint32_t)P.tile.w, pix / (uint32_t)P.tile.w, sidx,
                 &samid, &pixid, &d);
      float4* out = reinterpret_cast<float4*>(rays + i);
      out[0] = make_float4(o.x, o.y, o.z, SPRAY_RAY_EPS);
      out[1] = make_float4(d.x, d.y, d.z, INFINITY);
      mask = route_mask(boxes, R, P.nranks, o, d, &primary);
    }
    fwd_push(P, P.o_inbox_next, P.inbox_next, mask, primary, o, d, (uint32_t)i, (uint32_t)P.me);
  }
}

// F2: trace MY inbox against MY domains (insitu::TContext::processRays for received rays), in two passes separated by a
// barrier.  Pass 0: the records for which I own the NEAREST entered domain.  Pass 1: the others — by then the home
// rank's key / occlusion byte holds what the nearest owner found, so a radiance ray is traced only in front of that
// hit (tbound) and an already occluded shadow ray not at all.  In an opaque scene most rays end in the first domain
// they enter: without the second pass every other owner along the ray walks its BVHs for nothing (measured at 2 GPUs
// on C3: the speculative single pass cost 2x the single-GPU traversal work).  Persistent CTAs; the records of a pass
// are compacted through a per-warp index queue so that the traversal runs with full warps.
__global__ void __launch_bounds__(FWD_THREADS, FWD_MINBLOCKS)
    fwd_serve_kernel(FwdWave P, int pass, uint32_t* __restrict__ cand_idx, unsigned long long* __restrict__ cand_key) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(P.table, P.nd, P.tlas, P.tlas_n, smem);
  uint32_t* q = (uint32_t*)((char*)smem + ((dom_smem_bytes(P.nd, P.tlas_n, 0) + 15) & ~(size_t)15)) + (size_t)(threadIdx.x >> 5) * FWD_QCAP;
  const int lane = threadIdx.x & 31;
  uint32_t seg_end[P2P_MAX_RANKS];
  uint32_t nrec = 0;
#pragma unroll
  for (int r = 0; r < P2P_MAX_RANKS; ++r) {
    if (r < P.nranks) nrec += ((const unsigned int*)(P.peers.base[P.me] + FWD_CTR_OFFSET))[P.inbox_cur * P2P_MAX_RANKS + r];
    seg_end[r] = nrec;
  }
  const uint4* inbox = (const uint4*)(P.peers.base[P.me] + P.o_inbox_cur);
  unsigned int* work = P.ctl + (pass ? FC_WORK_ROUTE : FC_WORK_SERVE);   // the route counter is idle during a wave's F2
  int count = 0;
  bool more = true;
  for (;;) {
    if (count < 32 && more) {
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(work, 32u);
      base = __shfl_sync(0xffffffffu, base, 0);
      if (base >= nrec) { more = false; continue; }
      uint32_t k = base + lane;
      bool go = false;
      if (k < nrec) {
        k = fwd_phys(seg_end, P.nranks, P.seg_cap, k);
        const uint32_t meta = inbox[2 * (size_t)k + 1].w;
        go = (int)((meta >> 9) & 1u) == pass;
        if (go && pass && (meta & FWD_META_SHADOW)) {
          // VBuf::occluded: somebody nearer already blocked it
          const uint32_t sample = inbox[2 * (size_t)k].w;
          go = ((const volatile uint8_t*)(P.peers.base[meta & 0xffu] + P.o_occ))[(size_t)(meta >> 16) * P.n + sample] == 0;
        }
      }
      const uint32_t bm = __ballot_sync(0xffffffffu, go);
      if (go) q[count + __popc(bm & ((1u << lane) - 1u))] = k;
      count += __popc(bm);
      __syncwarp();
      continue;
    }
    if (count == 0) break;
    const int nq = count < 32 ? count : 32;
    count -= nq;
    const bool act = lane < nq;
    const uint32_t k = q[count + (act ? lane : 0)];
    __syncwarp();
    const uint4 a = inbox[2 * (size_t)k], b = inbox[2 * (size_t)k + 1];
    const vec3 o = mk3(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z));
    const vec3 d = mk3(__uint_as_float(b.x), __uint_as_float(b.y), __uint_as_float(b.z));
    const uint32_t sample = a.w, home = b.w & 0xffu, lpass = b.w >> 16;
    const bool shadow = (b.w & FWD_META_SHADOW) != 0u;
    unsigned long long* hkey = (unsigned long long*)(P.peers.base[home] + P.o_keys) + sample;
    float tbound = INFINITY;
    if (act && pass && !shadow) {
      const unsigned long long known = *(const volatile unsigned long long*)hkey;
      if (known != SPRAY_KEY_MISS) tbound = key_t(known);                                 // VBuf::correct: only nearer hits matter
    }
    SceneHit h;
    const bool hit = scene_trace_warp<false>(P.table, D, act, o, d, shadow, h, nullptr, tbound);
    bool candidate = false;
    unsigned long long key = 0;
    if (act && hit) {
      if (shadow) {
        ((uint8_t*)(P.peers.base[home] + P.o_occ))[(size_t)lpass * P.n + sample] = 1;    // VBuf::setObuf on the home rank
      } else {
        key = make_key(h.t, h.dom, h.prim);
        const unsigned long long old = atomicMin_system(hkey, key);
        candidate = key < old;                                                            // VBuf::updateTbufOut
      }
    }
    const uint32_t slot = warp_append(P.ctl + FC_CAND, candidate);
    if (candidate) { cand_idx[slot] = k; cand_key[slot] = key; }
  }
  grid_signal_peers(P.ctl + FC_DONE, P.peers, P.nranks, P.me);
}

// F3: TriMeshBuffer::updateIntersection by the owner of the winning domain, for the candidates whose key survived the
// cross-rank min; 16 B (Ns.xyz, colour) into the home rank's attribute array.  (u, v) are re-derived from (ray, prim)
// with the S1 arithmetic on the same vertex values: bit-identical to the traversal's.
__global__ void __launch_bounds__(128) fwd_capture_kernel(FwdWave P, const uint32_t* __restrict__ cand_idx,
                                                          const unsigned long long* __restrict__ cand_key) {
  const uint32_t ncand = P.ctl[FC_CAND];
  const uint4* inbox = (const uint4*)(P.peers.base[P.me] + P.o_inbox_cur);
  for (size_t c = blockIdx.x * (size_t)blockDim.x + threadIdx.x; c < ncand; c += (size_t)gridDim.x * blockDim.x) {
    const uint32_t k = cand_idx[c];
    const unsigned long long key = cand_key[c];
    const uint4 a = inbox[2 * (size_t)k], b = inbox[2 * (size_t)k + 1];
    const uint32_t sample = a.w, home = b.w & 0xffu;
    const unsigned long long fin = *((const volatile unsigned long long*)(P.peers.base[home] + P.o_keys) + sample);
    if (fin != key) continue;                      // a nearer hit (another domain, maybe another rank) won
    const DomainEntry* e = P.table + key_domain(key);
    const vec3 o = mk3(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z));
    const vec3 d = mk3(__uint_as_float(b.x), __uint_as_float(b.y), __uint_as_float(b.z));
    const uint32_t prim = key_prim(key);
    const uint32_t i0 = e->faces[3 * (size_t)prim], i1 = e->faces[3 * (size_t)prim + 1], i2 = e->faces[3 * (size_t)prim + 2];
    const vec3 v0 = mk3(e->verts[3 * (size_t)i0], e->verts[3 * (size_t)i0 + 1], e->verts[3 * (size_t)i0 + 2]);
    const vec3 v1 = mk3(e->verts[3 * (size_t)i1], e->verts[3 * (size_t)i1 + 1], e->verts[3 * (size_t)i1 + 2]);
    const vec3 v2 = mk3(e->verts[3 * (size_t)i2], e->verts[3 * (size_t)i2 + 1], e->verts[3 * (size_t)i2 + 2]);
    TriHit th; th.t = th.u = th.v = 0.f;
    tri_test(o, d, v0, gsub(v0, v1), gsub(v2, v0), SPRAY_RAY_EPS, &th);
    uint32_t col; vec3 Ns;
    fetch_attributes(e, prim, th.u, th.v, &col, &Ns);
    ((uint4*)(P.peers.base[home] + P.o_cap))[sample] = make_uint4(__float_as_uint(Ns.x), __float_as_uint(Ns.y), __float_as_uint(Ns.z), col);
  }
  grid_signal_peers(P.ctl + FC_DONE, P.peers, P.nranks, P.me);
}

// F4: the HOME rank shades its samples from (final key, attributes) — ShaderPt / ShaderAo (insitu_shader_pt.h:83-230,
// insitu_shader_ao.h) — once, not once per rank.  Shadow rays and the bounce ray are routed into the NEXT wave's
// inboxes right here; the bounce ray is also kept in rays_next / w_next / the live list for the next shading pass.
// Persistent CTAs; hit samples are compacted through a per-warp index queue so that shading runs with full warps.
__global__ void __launch_bounds__(FWD_THREADS, FWD_MINBLOCKS)
    fwd_shade_kernel(FwdWave P, const spraygpu_ray* __restrict__ rays, const float* __restrict__ w, spraygpu_ray* __restrict__ rays_next,
                     float* __restrict__ w_next, uint32_t* __restrict__ next_list, unsigned int* __restrict__ next_n,
                     uint4* __restrict__ ret, unsigned long long* __restrict__ counters) {
  extern __shared__ float smem[];
  float* boxes = smem;
  for (int i = threadIdx.x; i < 6 * P.nd; i += blockDim.x) boxes[i] = P.table[i / 6].box[i % 6];
  const RouteTab R = load_route_smem(P.route, P.nranks, P.nd, smem + 6 * P.nd);   // syncs
  uint32_t* q = (uint32_t*)(smem + 6 * P.nd + route_words(P.nranks, P.nd)) + (size_t)(threadIdx.x >> 5) * FWD_QCAP;
  const int lane = threadIdx.x & 31;
  const uint32_t nsrc = P.src ? *P.src_n : (uint32_t)P.n;
  unsigned long long* keys = (unsigned long long*)(P.peers.base[P.me] + P.o_keys);
  unsigned long long* keys_next = (unsigned long long*)(P.peers.base[P.me] + P.o_keys_next);
  const uint4* cap = (const uint4*)(P.peers.base[P.me] + P.o_cap);
  uint8_t* occ = (uint8_t*)(P.peers.base[P.me] + P.o_occ);
  const bool pt = P.shade.shader == SPRAYGPU_SHADER_PT;
  const int nl = pt ? P.shade.nlights : 1;
  const int next_depth = P.depth + 1;
  uint32_t n_hit = 0, n_shadow = 0, n_rad = 0;
  int count = 0;
  bool more_src = true;
  for (;;) {
    if (count < 32 && more_src) {
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(P.ctl + FC_WORK_SHADE, 32u);
      base = __shfl_sync(0xffffffffu, base, 0);
      if (base >= nsrc) { more_src = false; continue; }
      const uint32_t k = base + lane;
      bool go = false;
      uint32_t i = 0;
      if (k < nsrc) {
        i = P.src ? P.src[k] : k;
        if (!P.src) { int row; fwd_pixid(P, i, &row); go = fwd_home(P, row) == P.me; } else go = true;
        go = go && keys[i] != SPRAY_KEY_MISS;
      }
      const uint32_t b = __ballot_sync(0xffffffffu, go);
      if (go) q[count + __popc(b & ((1u << lane) - 1u))] = i;
      count += __popc(b);
      __syncwarp();
      continue;
    }
    if (count == 0) break;
    const int nq = count < 32 ? count : 32;
    count -= nq;
    const bool hit = lane < nq;
    const uint32_t i = q[count + (hit ? lane : 0)];
    __syncwarp();
    const unsigned long long key = keys[i];
    const spraygpu_ray ry = load_ray(rays + i);
    const vec3 ro = mk3(ry.ox, ry.oy, ry.oz), rd = mk3(ry.dx, ry.dy, ry.dz);
    int row = 0;
    const uint32_t pixid = fwd_pixid(P, i, &row);
    const uint32_t samid = i;
    HitCtx c;
    c.pos = ro; c.kd = mk3(0, 0, 0); c.nff = mk3(0, 0, 1); c.wo = gneg(rd);
    c.cos_theta_i = 0.f; c.entering = false; c.mtl = SPRAYGPU_MTL_DIFFUSE; c.eta0 = c.eta1 = 1.f;
    vec3 Lin = mk3(0, 0, 0), bd = mk3(0, 0, 1), bw = mk3(0, 0, 0);
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
      Lin = P.depth ? mk3(w[3 * (size_t)i], w[3 * (size_t)i + 1], w[3 * (size_t)i + 2]) : mk3(1.0f, 1.0f, 1.0f);
      ++n_hit;
      if (next_depth < P.shade.bounces) {
        bool em1, em2; vec3 d1 = mk3(0, 0, 1), w1 = mk3(0, 0, 0), d2 = mk3(0, 0, 1), w2 = mk3(0, 0, 0);
        shade_secondary<true>(c, Lin, samid, next_depth, &em1, &d1, &w1, &em2, &d2, &w2);
        // glass domains are rejected at launch, so at most one of the two is live
        if (em1) { bd = d1; bw = w1; bounce = true; }
        else if (em2) { bd = d2; bw = w2; bounce = true; }
        if (bounce) ++n_rad;
      }
    }
    // the bounce ray: next wave's radiance ray of this sample
    {
      const uint32_t slot = warp_append(next_n, bounce);
      uint32_t mask = 0;
      int primary = -1;
      if (bounce) {
        next_list[slot] = i;
        spraygpu_ray nx;
        nx.ox = c.pos.x; nx.oy = c.pos.y; nx.oz = c.pos.z; nx.tmin = SPRAY_RAY_EPS;
        nx.dx = bd.x; nx.dy = bd.y; nx.dz = bd.z; nx.tmax = INFINITY;
        store_ray(rays_next + i, nx);
        w_next[3 * (size_t)i] = bw.x; w_next[3 * (size_t)i + 1] = bw.y; w_next[3 * (size_t)i + 2] = bw.z;
        keys_next[i] = SPRAY_KEY_MISS;                          // VBuf::resetTbufOut for the samples that go on
        mask = route_mask(boxes, R, P.nranks, c.pos, bd, &primary);
      }
      fwd_push(P, P.o_inbox_next, P.inbox_next, mask, primary, c.pos, bd, i, (uint32_t)P.me);
    }
    uint32_t ls = rng_init1(samid * (uint32_t)next_depth);     // light_sampler (insitu_shader_pt.h:115)
    int l = 0, k = 0;
    bool more = hit && c.mtl == SPRAYGPU_MTL_DIFFUSE && nl > 0;
    for (int p = 0; p < P.npasses; ++p) {
      if (!__any_sync(0xffffffffu, more)) break;
      bool emit = false; vec3 wi = mk3(0, 0, 1), Lr = mk3(0, 0, 0);
      if (more) emit = shade_light_pass<true>(P.shade, pt, nl, c, Lin, pixid, l, k, ls, &wi, &Lr, &more);
      const size_t j = (size_t)p * P.n + i;
      uint32_t mask = 0;
      int primary = -1;
      if (emit) {
        ++n_shadow;
        occ[j] = 0;
        mask = route_mask(boxes, R, P.nranks, c.pos, wi, &primary);
      }
      fwd_push(P, P.o_inbox_next, P.inbox_next, mask, primary, c.pos, wi, i, (uint32_t)P.me | FWD_META_SHADOW | ((uint32_t)p << 16));
      const uint32_t slot = warp_append(P.ctl + FC_RET, emit);
      if (emit) ret[slot] = make_uint4((uint32_t)j, __float_as_uint(Lr.x), __float_as_uint(Lr.y), __float_as_uint(Lr.z));
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
}

// F5: TContext::retireShadows -> HdrImage::add (all entries are this rank's home samples)
__global__ void __launch_bounds__(256) fwd_retire_kernel(FwdWave P, const uint4* __restrict__ ret, float* __restrict__ fb) {
  const unsigned int nret = P.ctl[FC_RET];
  const uint8_t* occ = (const uint8_t*)(P.peers.base[P.me] + P.o_occ);
  const float scale = (float)(1.0 / (double)P.spp);
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < nret; e += (size_t)gridDim.x * blockDim.x) {
    const uint4 r = ret[e];
    if (((const volatile uint8_t*)occ)[r.x]) continue;
    int row;
    const uint32_t pixid = fwd_pixid(P, (size_t)r.x % P.n, &row);
    fb_add(fb, pixid, scale, mk3(__uint_as_float(r.y), __uint_as_float(r.z), __uint_as_float(r.w)));
  }
}

}  // namespace spraygpu

using namespace spraygpu;

static void fwd_free_local(FwdState* S) {
  for (int k = 0; k < 2; ++k) {
    if (S->rays[k]) cudaFree(S->rays[k]);
    if (S->w[k]) cudaFree(S->w[k]);
    if (S->list[k]) cudaFree(S->list[k]);
    S->rays[k] = nullptr; S->w[k] = nullptr; S->list[k] = nullptr;
  }
  if (S->ret) cudaFree(S->ret);
  if (S->cand_idx) cudaFree(S->cand_idx);
  if (S->cand_key) cudaFree(S->cand_key);
  S->ret = nullptr; S->cand_idx = nullptr; S->cand_key = nullptr;
}
static void fwd_destroy(Comm* c, void* p) {
  FwdState* S = (FwdState*)p;
  if (!S) return;
  cudaSetDevice(comm_device(c)->cuda_device);
  cudaStreamSynchronize(comm_device(c)->stream);
  if (S->local) { comm_host_barrier(c); comm_shared_free(c, S->local, &S->mapped); }
  fwd_free_local(S);
  if (S->d_ctl) cudaFree(S->d_ctl);
  if (S->d_route) cudaFree(S->d_route);
  if (S->h_err) cudaFreeHost(S->h_err);
  delete S;
}

// collective: every rank calls with the same (n, npasses)
static int fwd_ensure(Comm* c, FwdState* S, size_t n, int npasses, size_t need_seg) {
  Device* dev = comm_device(c);
  if (!S->d_ctl) {
    SG_CUDA(cudaMalloc((void**)&S->d_ctl, sizeof(unsigned int) * FC_WORDS));
    SG_CUDA(cudaMemset(S->d_ctl, 0, sizeof(unsigned int) * FC_WORDS));
    SG_CUDA(cudaMalloc((void**)&S->d_route, sizeof(float) * route_words(P2P_MAX_RANKS, SPRAYGPU_MAX_DOMAINS)));
    SG_CUDA(cudaHostAlloc((void**)&S->h_err, sizeof(unsigned int), cudaHostAllocMapped));
    *S->h_err = 0u;
    SG_CUDA(cudaHostGetDevicePointer((void**)&S->h_err_dev, S->h_err, 0));
  }
  if (n <= S->cap_n && npasses <= S->cap_passes && need_seg <= S->seg_cap) return SPRAYGPU_OK;
  SG_CUDA(cudaStreamSynchronize(dev->stream));
  if (S->local) {
    int rc = comm_host_barrier(c);
    if (rc) return rc;
    comm_shared_free(c, S->local, &S->mapped);
    S->local = nullptr;
  }
  fwd_free_local(S);
  const size_t cn = std::max(n, S->cap_n);
  const int cp = std::max(npasses, std::max(S->cap_passes, 1));
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  // worst case: every home sample of a source rank sends its radiance ray and all its shadow rays to this rank
  S->seg_cap = std::max(need_seg, S->seg_cap);
  const size_t inbox_cap = S->seg_cap * (size_t)comm_size(c);
  size_t off = P2P_FLAG_BYTES;
  for (int k = 0; k < 2; ++k) { S->o_inbox[k] = off; off += al(inbox_cap * 32); }
  for (int k = 0; k < 2; ++k) { S->o_keys[k] = off; off += al(cn * 8); }
  S->o_cap = off; off += al(cn * 16);
  S->o_occ = off; off += al(cn * (size_t)cp);
  int rc = comm_shared_alloc(c, off, &S->local, &S->mapped, -1);   // zero-filled: flag = 0, inbox counters = 0
  if (rc) return rc;
  S->seq = 0;
  for (int k = 0; k < 2; ++k) {
    SG_CUDA(cudaMalloc((void**)&S->rays[k], cn * sizeof(spraygpu_ray)));
    SG_CUDA(cudaMalloc((void**)&S->w[k], cn * 12));
    SG_CUDA(cudaMalloc((void**)&S->list[k], cn * sizeof(uint32_t)));
  }
  SG_CUDA(cudaMalloc((void**)&S->ret, cn * sizeof(uint4) * (size_t)cp));
  SG_CUDA(cudaMalloc((void**)&S->cand_idx, inbox_cap * sizeof(uint32_t)));
  SG_CUDA(cudaMalloc((void**)&S->cand_key, inbox_cap * sizeof(unsigned long long)));
  S->cap_n = cn; S->cap_passes = cp;
  return comm_host_barrier(c);
}

extern "C" int spraygpu_trace_tile_insitu_fwd(spraygpu_device dev_h, spraygpu_comm comm_h, const spraygpu_camera* cam,
                                              const spraygpu_shade* shade, spraygpu_tile tile, float* fb, int rows_per_block,
                                              const int32_t* domain_owner, spraygpu_stats* stats) {
  SG_CHECK(dev_h && comm_h && cam && shade && fb && domain_owner, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: NULL argument");
  Comm* comm = (Comm*)comm_h;
  Device* dev = (Device*)dev_h;
  SG_CHECK(comm_device(comm) == dev, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: communicator belongs to another device");
  const int nranks = comm_size(comm), me = comm_rank(comm);
  SG_CHECK(nranks <= P2P_MAX_RANKS, SPRAYGPU_ERR_LIMIT, "trace_tile_insitu_fwd: %d ranks (limit %d)", nranks, P2P_MAX_RANKS);
  SG_CHECK(rows_per_block >= 1, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: rows_per_block must be >= 1");
  SG_CHECK(shade->pixel_samples >= 1 && shade->bounces >= 1, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: pixel_samples and bounces must be >= 1");
  SG_CHECK(shade->nlights >= 0 && shade->nlights <= SPRAYGPU_MAX_LIGHTS, SPRAYGPU_ERR_LIMIT, "trace_tile_insitu_fwd: %d lights", shade->nlights);
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_AO || shade->nlights > 0, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: the PT shader needs a light");
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_PT || shade->ao_samples > 0, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: ao_samples must be > 0");
  SG_CHECK(tile.x >= 0 && tile.y >= 0 && tile.w >= 0 && tile.h >= 0 && tile.x + tile.w <= cam->image_w &&
               tile.y + tile.h <= cam->image_h, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: tile outside the image");
  SG_CHECK(dev->ndomains > 0, SPRAYGPU_ERR_STATE, "trace_tile_insitu_fwd: no domain table (spraygpu_domains_set)");
  SG_CUDA(cudaSetDevice(dev->cuda_device));
  {
    cudaPointerAttributes a;
    const bool devptr = cudaPointerGetAttributes(&a, fb) == cudaSuccess && (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    SG_CHECK(devptr, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: fb must be a device pointer");
  }
  const int nd = dev->ndomains;
  for (int i = 0; i < nd; ++i) {
    SG_CHECK(dev->h_domains[i].mtl_type != SPRAYGPU_MTL_GLASS, SPRAYGPU_ERR_LIMIT,
             "trace_tile_insitu_fwd: glass domains (two secondary rays per hit) are not supported on the in-situ path; "
             "the reference CHECK-fails on them too (insitu_shader_pt.h)");
    SG_CHECK(domain_owner[i] >= 0 && domain_owner[i] < nranks, SPRAYGPU_ERR_ARG, "trace_tile_insitu_fwd: domain %d has owner %d of %d ranks", i,
             domain_owner[i], nranks);
    SG_CHECK((domain_owner[i] == me) == (dev->h_domains[i].resident != 0), SPRAYGPU_ERR_STATE,
             "trace_tile_insitu_fwd: domain %d: owner %d but resident = %d on rank %d (a rank must hold exactly its own domains)", i,
             domain_owner[i], dev->h_domains[i].resident, me);
  }
  void (**free_slot)(Comm*, void*) = nullptr;
  void** slot = comm_p2p_slot(comm, &free_slot, 1);
  if (!*slot) { *slot = new FwdState(); *free_slot = fwd_destroy; }
  FwdState* S = (FwdState*)*slot;
  { int erc = p2p_take_error(S->h_err, "trace_tile_insitu_fwd"); if (erc) return erc; }
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
    SG_CHECK(n * (size_t)(npasses + 1) < ((size_t)1 << 32), SPRAYGPU_ERR_LIMIT,
             "trace_tile_insitu_fwd: %zu samples x %d light passes do not fit 32-bit record indices; use smaller blocking tiles", n, npasses);
    SG_CHECK(npasses < 65536, SPRAYGPU_ERR_LIMIT, "trace_tile_insitu_fwd: %d light passes", npasses);
    // a rank is home to at most ceil(h / (rows_per_block * nranks)) row blocks of the tile
    const size_t home_rows = (size_t)((tile.h + rows_per_block * nranks - 1) / (rows_per_block * nranks) + 1) * rows_per_block;
    const size_t need_seg = std::min(home_rows, (size_t)tile.h) * tile.w * shade->pixel_samples * (size_t)(1 + npasses);
    SG_CHECK(need_seg * (size_t)nranks < ((size_t)1 << 32), SPRAYGPU_ERR_LIMIT,
             "trace_tile_insitu_fwd: %zu inbox records do not fit 32-bit record indices; use smaller blocking tiles", need_seg * (size_t)nranks);
    rc = fwd_ensure(comm, S, n, npasses, need_seg);
    if (rc) return rc;
    // routing table: domains grouped by owner, one box per rank
    {
      std::vector<float> rt(route_words(nranks, nd));
      float* rbox = rt.data();
      int* begin = (int*)(rt.data() + 6 * nranks);
      int* perm = begin + nranks + 1;
      int k = 0;
      for (int r = 0; r < nranks; ++r) {
        begin[r] = k;
        float* b = rbox + 6 * r;
        b[0] = b[1] = b[2] = INFINITY; b[3] = b[4] = b[5] = -INFINITY;
        for (int i = 0; i < nd; ++i)
          if (domain_owner[i] == r) {
            perm[k++] = i;
            for (int a = 0; a < 3; ++a) {
              b[a] = std::min(b[a], dev->h_domains[i].box[a]);
              b[3 + a] = std::max(b[3 + a], dev->h_domains[i].box[3 + a]);
            }
          }
      }
      begin[nranks] = k;
      SG_CUDA(cudaMemcpyAsync(S->d_route, rt.data(), sizeof(float) * rt.size(), cudaMemcpyHostToDevice, st));   // pageable: staged before return
    }
    FwdWave P;
    P.table = dev->d_domains; P.nd = nd; P.tlas = dev->d_tlas; P.tlas_n = dev->tlas_n; P.route = S->d_route;
    P.shade = *shade; P.tile = tile; P.cam = to_camdev(cam); P.image_w = cam->image_w; P.spp = shade->pixel_samples;
    P.npasses = npasses; P.n = n; P.nranks = nranks; P.me = me; P.rows_per_block = rows_per_block;
    P.o_cap = S->o_cap; P.o_occ = S->o_occ; P.ctl = S->d_ctl; P.seg_cap = (uint32_t)S->seg_cap;
    // SPRAY_FWD_SEG_CAP (test knob, the same on every rank): pretend the inbox segments are this small, so the overflow
    // report can be exercised — need_seg above is a worst-case bound that real tiles never exceed
    if (const char* e = getenv("SPRAY_FWD_SEG_CAP")) { const long v = atol(e); if (v > 0 && (size_t)v < S->seg_cap) P.seg_cap = (uint32_t)v; }
    for (int r = 0; r < P2P_MAX_RANKS; ++r) P.peers.base[r] = r < nranks ? (char*)S->mapped[(size_t)r] : nullptr;
    char* mine = (char*)S->local;
    auto barrier_wait = [&]() {
      ++S->seq;
      p2p_wait_kernel<<<1, 1, 0, st>>>((volatile unsigned long long*)mine, S->seq * (unsigned long long)(nranks - 1), S->h_err_dev, p2p_timeout_cycles());
      dev->launches++;
    };
    const size_t dsmem = (dom_smem_bytes(nd, dev->tlas_n, 0) + 15) & ~(size_t)15;
    const size_t smem_route = sizeof(float) * (6 * (size_t)nd + route_words(nranks, nd));
    const size_t smem_shade = smem_route + sizeof(uint32_t) * FWD_QCAP * (FWD_THREADS / 32);
    const size_t smem_serve = dsmem + sizeof(uint32_t) * FWD_QCAP * (FWD_THREADS / 32);
    if (S->grid_smem != dsmem + smem_shade) {
      SG_DYN_SMEM(fwd_serve_kernel, smem_serve);
      SG_DYN_SMEM(fwd_shade_kernel, smem_shade);
      SG_DYN_SMEM(fwd_route0_kernel, smem_route);
      int b2 = 0, b4 = 0;
      SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, fwd_serve_kernel, FWD_THREADS, smem_serve));
      SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b4, fwd_shade_kernel, FWD_THREADS, smem_shade));
      SG_CHECK(b2 > 0 && b4 > 0, SPRAYGPU_ERR_LIMIT, "trace_tile_insitu_fwd: the wave kernels do not fit an SM (%zu B shared)", dsmem);
      S->grid_serve = b2 * dev->sm_count; S->grid_shade = b4 * dev->sm_count; S->grid_smem = dsmem + smem_shade;
    }
    const int grid_lin = dev->sm_count * 8;
    PhaseTimer T(st);
    // ---- tile start: my wave-0 keys reset, both inboxes empty, then everyone meets
    SG_CUDA(cudaMemsetAsync(mine + S->o_keys[0], 0xFF, n * 8, st));
    fwd_reset_kernel<<<1, 1, 0, st>>>(S->d_ctl, 3, 0);
    fwd_zero_ret_kernel<<<1, 1, 0, st>>>(S->d_ctl);
    p2p_signal_kernel<<<1, 1, 0, st>>>(P.peers, nranks, me);
    dev->launches += 3;
    barrier_wait();
    T.mark("init+bar");
    // ---- F1: eye rays into inbox 0
    P.depth = 0; P.src = nullptr; P.src_n = S->d_ctl + FC_LIVE0;
    P.inbox_next = 0; P.o_inbox_next = S->o_inbox[0]; P.inbox_cur = 1; P.o_inbox_cur = S->o_inbox[1];
    P.o_keys = S->o_keys[0]; P.o_keys_next = S->o_keys[1];
    fwd_route0_kernel<<<grid_lin, FWD_THREADS, smem_route, st>>>(P, S->rays[0]);
    fwd_publish_kernel<<<1, 1, 0, st>>>(P.peers, nranks, me, S->d_ctl, 0, P.seg_cap);
    dev->launches += 2;
    T.mark("route0");
    barrier_wait();
    T.mark("bar");
    int cur = 0;
    const int nwaves = shade->bounces + (npasses ? 1 : 0);     // the last one serves the trailing shadow rays only
    for (int depth = 0; depth < nwaves; ++depth) {
      const int kb = depth & 1;
      const bool shading = depth < shade->bounces;
      P.depth = depth;
      P.inbox_cur = kb; P.o_inbox_cur = S->o_inbox[kb]; P.inbox_next = kb ^ 1; P.o_inbox_next = S->o_inbox[kb ^ 1];
      P.o_keys = S->o_keys[kb]; P.o_keys_next = S->o_keys[kb ^ 1];
      P.src = depth ? S->list[kb] : nullptr;
      P.src_n = S->d_ctl + FC_LIVE0 + kb;
      // nobody appends to the next inbox before the barrier that follows F3 of this wave
      fwd_reset_kernel<<<1, 1, 0, st>>>(S->d_ctl, 1 << (kb ^ 1), kb ^ 1);
      fwd_serve_kernel<<<S->grid_serve, FWD_THREADS, smem_serve, st>>>(P, 0, S->cand_idx, S->cand_key);       // F2, nearest owners
      dev->launches += 2;
      T.mark("serve");
      barrier_wait();
      T.mark("bar");
      if (nranks > 1) {
        fwd_serve_kernel<<<S->grid_serve, FWD_THREADS, smem_serve, st>>>(P, 1, S->cand_idx, S->cand_key);     // F2, the others, bounded
        dev->launches++;
        T.mark("serve2");
        barrier_wait();
        T.mark("bar");
      }
      if (depth > 0 && npasses) {
        fwd_retire_kernel<<<grid_lin, 256, 0, st>>>(P, S->ret, fb);                                              // F5 (wave depth - 1)
        fwd_zero_ret_kernel<<<1, 1, 0, st>>>(S->d_ctl);
        dev->launches += 2;
        T.mark("retire");
      }
      if (!shading) break;
      fwd_capture_kernel<<<grid_lin, 128, 0, st>>>(P, S->cand_idx, S->cand_key);                                // F3
      dev->launches++;
      T.mark("capture");
      barrier_wait();
      T.mark("bar");
      fwd_shade_kernel<<<S->grid_shade, FWD_THREADS, smem_shade, st>>>(P, S->rays[cur], S->w[cur], S->rays[cur ^ 1], S->w[cur ^ 1],
                                                                       S->list[kb ^ 1], S->d_ctl + FC_LIVE0 + (kb ^ 1), S->ret,
                                                                       dev->d_counters);                        // F4
      fwd_publish_kernel<<<1, 1, 0, st>>>(P.peers, nranks, me, S->d_ctl, kb ^ 1, P.seg_cap);
      dev->launches += 2;
      T.mark("shade");
      barrier_wait();
      T.mark("bar");
      cur ^= 1;
    }
    fwd_report_kernel<<<1, 1, 0, st>>>(S->d_ctl, S->h_err_dev);
    dev->launches++;
    T.report("insitu_fwd", me);
    SG_CUDA(cudaGetLastError());
  }
  if (stats) {
    rc = comm_allreduce_u64_sum(comm, dev->d_counters, 8);      // every rank counted its home samples
    if (rc) return rc;
    unsigned long long* h = (unsigned long long*)dev->pinned;
    SG_CUDA(cudaMemcpyAsync(h, dev->d_counters, sizeof(unsigned long long) * 8, cudaMemcpyDeviceToHost, st));
    SG_CUDA(cudaStreamSynchronize(st));
    { int erc = p2p_take_error(S->h_err, "trace_tile_insitu_fwd"); if (erc) return erc; }   // this tile's barriers and inboxes
    stats->eye_rays = n; stats->radiance_rays = n + h[2]; stats->shadow_rays = h[3]; stats->domain_tests = 0;
    stats->nodes_visited = 0; stats->tris_tested = 0; stats->hits = h[7];
  }
  return SPRAYGPU_OK;
}
