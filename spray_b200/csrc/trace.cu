// trace.cu — ray kernels of libspraygpu.
//
//   render_tile_kernel      the fused persistent kernel: eye ray -> domain culling -> per-domain BVH8
//                           closest hit -> shading -> shadow rays (any hit) -> bounce -> frame buffer.
//                           One lane = one sample's whole ray stream; warps pull 32 samples at a time
//                           from a global counter (8x4 pixel blocks for coherence).
//   intersect_domains_kernel / occluded_domains_kernel   batched queries over all resident domains
//                           (64-bit atomicMin merge for multi-GPU / out-of-core accumulation)
//   intersect_scene_kernel / occluded_scene_kernel       batched single-domain queries (seam 2)
//   domain_lists_kernel, hit_attributes_kernel, eye_rays_kernel   parity surfaces for a6 / a3 / a9
#include <stdlib.h>

#include "internal.cuh"
#include "traverse.cuh"
#include "shade.cuh"

namespace spraygpu {

// ------------------------------------------------------------------------------------------------
// batched kernels.  No thread leaves before the warp-cooperative culling: out-of-range lanes stay as
// inactive participants.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
    domain_lists_kernel(const DomainEntry* __restrict__ table, int nd, const TlasNode* __restrict__ tlas, int tlas_n,
                        const spraygpu_ray* __restrict__ rays, size_t n, int cap, int32_t* __restrict__ counts, int32_t* __restrict__ ids, float* __restrict__ tmins) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(table, nd, tlas, tlas_n, smem);
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool valid = i < n;
  spraygpu_ray ry = {0, 0, 0, 0, 0, 0, 1, 0};
  if (valid) ry = load_ray(rays + (i));
  const vec3 o = mk3(ry.ox, ry.oy, ry.oz), d = mk3(ry.dx, ry.dy, ry.dz);
  warp_cull(D, valid, o, d, false);
  if (!valid) return;
  uint64_t list[MAXLIST];
  bool have_after = false, more = false;
  uint64_t after = 0;
  int total = 0;
  do {
    const int c = domain_list(D, WarpCand{D.cand}, o, d, have_after, after, list, &more);
    for (int k = 0; k < c; ++k, ++total) {
      if (total < cap) {
        ids[i * cap + total] = (int32_t)(uint32_t)list[k];
        tmins[i * cap + total] = sortable_float((uint32_t)(list[k] >> 32));
      }
    }
    if (c > 0) { after = list[c - 1]; have_after = true; }
  } while (more);
  counts[i] = total;
}

__global__ void __launch_bounds__(128)
    intersect_scene_kernel(const uint4* __restrict__ nodes, const float4* __restrict__ tris, uint32_t domain_id,
                           const spraygpu_ray* __restrict__ rays, size_t n, uint64_t* __restrict__ keys,
                           float* __restrict__ uv) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool valid = i < n;
  spraygpu_ray ry = {0, 0, 0, 0, 0, 0, 1, 0};
  if (valid) ry = load_ray(rays + (i));
  const RayCtx r = make_ray_ctx(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), ry.tmin);
  OneDomainSource src{nodes, tris, domain_id, false};
  SceneHit c;
  const bool hit = trace_warp<false>(src, valid, r, false, c, nullptr);
  if (!valid) return;
  keys[i] = hit ? make_key(c.t, domain_id, c.prim) : SPRAY_KEY_MISS;
  if (uv) { uv[2 * i] = hit ? c.u : 0.f; uv[2 * i + 1] = hit ? c.v : 0.f; }
}

__global__ void __launch_bounds__(128)
    occluded_scene_kernel(const uint4* __restrict__ nodes, const float4* __restrict__ tris,
                          const spraygpu_ray* __restrict__ rays, size_t n, uint8_t* __restrict__ occ) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool valid = i < n;
  spraygpu_ray ry = {0, 0, 0, 0, 0, 0, 1, 0};
  if (valid) ry = load_ray(rays + (i));
  const RayCtx r = make_ray_ctx(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), ry.tmin);
  OneDomainSource src{nodes, tris, 0u, false};
  SceneHit c;
  const bool oc = trace_warp<false>(src, valid, r, true, c, nullptr);
  if (valid) occ[i] = oc ? 1 : 0;
}

__global__ void __launch_bounds__(128)
    intersect_domains_kernel(const DomainEntry* __restrict__ table, int nd, const TlasNode* __restrict__ tlas, int tlas_n,
                             const spraygpu_ray* __restrict__ rays, size_t n, uint64_t* __restrict__ keys, float* __restrict__ uv, int merge) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(table, nd, tlas, tlas_n, smem);
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool valid = i < n;
  spraygpu_ray ry = {0, 0, 0, 0, 0, 0, 1, 0};
  if (valid) ry = load_ray(rays + (i));
  const vec3 o = mk3(ry.ox, ry.oy, ry.oz), d = mk3(ry.dx, ry.dy, ry.dz);
  // a zero direction marks an inactive slot of a ray wave (terminated sample): it hits nothing
  const bool live = valid && !(d.x == 0.0f && d.y == 0.0f && d.z == 0.0f);
  SceneHit h;
  const bool hit = scene_trace_warp<false>(table, D, live, o, d, false, h, nullptr);
  if (!valid) return;
  const uint64_t key = hit ? make_key(h.t, h.dom, h.prim) : SPRAY_KEY_MISS;
  if (merge) {
    // speculative resolution: whichever GPU / pass holds the smallest (t, domain, prim) wins
    const uint64_t old = atomicMin((unsigned long long*)&keys[i], (unsigned long long)key);
    if (uv && key < old) { uv[2 * i] = h.u; uv[2 * i + 1] = h.v; }
  } else {
    keys[i] = key;
    if (uv) { uv[2 * i] = hit ? h.u : 0.f; uv[2 * i + 1] = hit ? h.v : 0.f; }
  }
}

__global__ void __launch_bounds__(128)
    occluded_domains_kernel(const DomainEntry* __restrict__ table, int nd, const TlasNode* __restrict__ tlas, int tlas_n,
                            const spraygpu_ray* __restrict__ rays, size_t n, uint8_t* __restrict__ occ, int merge) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(table, nd, tlas, tlas_n, smem);
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  // merge: rays already occluded elsewhere (or flagged 1 by the caller as "not a ray") are skipped
  const bool valid = i < n && !(merge && occ[i]);
  spraygpu_ray ry = {0, 0, 0, 0, 0, 0, 1, 0};
  if (valid) ry = load_ray(rays + (i));
  const vec3 o = mk3(ry.ox, ry.oy, ry.oz), d = mk3(ry.dx, ry.dy, ry.dz);
  const bool live = valid && !(d.x == 0.0f && d.y == 0.0f && d.z == 0.0f);
  SceneHit hs;
  const bool oc = scene_trace_warp<false>(table, D, live, o, d, true, hs, nullptr);
  if (!valid) return;
  if (merge) { if (oc) occ[i] = 1; }
  else occ[i] = oc ? 1 : 0;
}

__global__ void __launch_bounds__(128)
    hit_attributes_kernel(const DomainEntry* __restrict__ table, const spraygpu_ray* __restrict__ rays,
                          const uint64_t* __restrict__ keys, size_t n, uint32_t* __restrict__ color,
                          float* __restrict__ ns3) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t key = keys[i];
  uint32_t col = 0; vec3 Ns = mk3(0, 0, 0);
  if (key != SPRAY_KEY_MISS) {
    const DomainEntry* e = table + key_domain(key);
    if (e->resident) {
      // re-derive (u,v) from (ray, primID) with the same S1 arithmetic on the same vertex values: the
      // result is bit-identical to what the traversal computed, so no (u,v) has to travel with the key.
      const spraygpu_ray ry = load_ray(rays + (i));
      const uint32_t prim = key_prim(key);
      const uint32_t i0 = e->faces[3 * (size_t)prim], i1 = e->faces[3 * (size_t)prim + 1], i2 = e->faces[3 * (size_t)prim + 2];
      const vec3 v0 = mk3(e->verts[3 * (size_t)i0], e->verts[3 * (size_t)i0 + 1], e->verts[3 * (size_t)i0 + 2]);
      const vec3 v1 = mk3(e->verts[3 * (size_t)i1], e->verts[3 * (size_t)i1 + 1], e->verts[3 * (size_t)i1 + 2]);
      const vec3 v2 = mk3(e->verts[3 * (size_t)i2], e->verts[3 * (size_t)i2 + 1], e->verts[3 * (size_t)i2 + 2]);
      TriHit th; th.t = th.u = th.v = 0.f;
      tri_test(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), v0, gsub(v0, v1), gsub(v2, v0), ry.tmin, &th);
      fetch_attributes(e, prim, th.u, th.v, &col, &Ns);
    }
  }
  color[i] = col;
  ns3[3 * i] = Ns.x; ns3[3 * i + 1] = Ns.y; ns3[3 * i + 2] = Ns.z;
}

__global__ void __launch_bounds__(256)
    eye_rays_kernel(CameraDev cam, spraygpu_tile tile, int spp, int eye_seed, spraygpu_ray* __restrict__ rays) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t n = (size_t)tile.w * tile.h * spp;
  if (i >= n) return;
  const uint32_t sidx = (uint32_t)(i % spp);
  const uint32_t pix = (uint32_t)(i / spp);
  uint32_t samid, pixid; vec3 d;
  eye_sample(cam, tile, tile.w, 0, 0, spp, eye_seed, pix % (uint32_t)tile.w, pix / (uint32_t)tile.w, sidx, &samid, &pixid, &d);
  // two 16-byte stores per ray (the 8 scalar stores of a struct copy ran at ~0.9 TB/s: ncu r01s)
  float4* out = reinterpret_cast<float4*>(rays + samid);
  out[0] = make_float4(cam.pos[0], cam.pos[1], cam.pos[2], SPRAY_RAY_EPS);
  out[1] = make_float4(d.x, d.y, d.z, INFINITY);
}

// ------------------------------------------------------------------------------------------------
// The fused persistent render kernel.
//
// Persistent CTAs (SM count x resident CTAs/SM); each WARP pulls 32 sample slots (one 8x4 pixel block) at a
// time from a global counter and carries them through the whole ray stream in warp-synchronous
// generations:   closest-hit rays -> shade -> shadow rays (one light sample per pass) -> bounce rays.
// Every stage starts with the warp-cooperative domain culling above, so the lanes of a warp stay in
// the same stage (their traversals still diverge in length, as in any per-lane BVH walk).
// Reference: ooc::Tracer::trace tile body (ooc_tracer.inl:184-234), TContext::procRads/procShads/retire
// (ooc_tcontext.inl:29-135), ShaderPt / ShaderAo (ooc_shader_pt.h:94-227, ooc_shader_ao.h:92-197).
// ------------------------------------------------------------------------------------------------
#define SPRAY_MAX_FRAME_TILES 32
// 8x4 pixel blocks a warp claims per atomicAdd on the frame's work counter (A/B knob: a same-address atomic costs ~2 ns
// of L2 time and a 2048^2 frame is 131 K blocks)
#ifndef SPRAY_BLOCK_SWIZZLE_DEFAULT
#define SPRAY_BLOCK_SWIZZLE_DEFAULT 0
#endif
#ifndef SPRAY_CLAIM_BLOCKS
#define SPRAY_CLAIM_BLOCKS 1
#endif
struct RenderParams {
  const DomainEntry* table;
  int nd;
  const TlasNode* tlas;
  int tlas_n;
  CameraDev cam;
  spraygpu_shade shade;
  float* fb;                     // image_w * image_h * 4
  unsigned long long* counters;  // [0] work counter; [1..7] stats
  // One launch walks up to SPRAY_MAX_FRAME_TILES blocking tiles (a whole frame of the reference's tile loop,
  // ooc_tracer.inl:184-234): the sample slots of the tiles are concatenated, every tile's share is a multiple of
  // 32, so a warp's 32 slots never straddle tiles and samid stays the index WITHIN the tile.
  int ntiles;
  spraygpu_tile tiles[SPRAY_MAX_FRAME_TILES];       // the rectangles rendered by this launch
  // position of each rectangle inside the blocking tile that numbers its samples (eye_sample): width of the blocking
  // tile and the rectangle's offset in it; (tile.w, 0, 0) when a whole blocking tile is rendered
  int32_t parent_w[SPRAY_MAX_FRAME_TILES];
  uint16_t parent_dx[SPRAY_MAX_FRAME_TILES], parent_dy[SPRAY_MAX_FRAME_TILES];
  uint32_t slot_end[SPRAY_MAX_FRAME_TILES];        // exclusive prefix end of tile t's slots (32 * pixel blocks * spp)
  uint32_t blocks_per_row[SPRAY_MAX_FRAME_TILES];
  // block enumeration inside a tile: 0 = row-major; else super-tiles of 2^(swz & 15) x 2^(swz >> 4) pixel blocks, row-major
  // inside a super-tile and across super-tiles (blocks_per_row is then padded to whole super-tiles): the warps that
  // are active at the same time cover a compact patch of the image instead of a band as wide as the tile
  uint8_t swz[SPRAY_MAX_FRAME_TILES];
  uint32_t total_slots;
  int qcap;                      // GENERAL variant: ray-queue entries per warp (the lean variant's is a constant)
};

__device__ __forceinline__ void block_xy(uint32_t blk, uint32_t bpr, uint32_t swz, uint32_t* bx, uint32_t* by) {
  if (swz == 0u) { *bx = blk % bpr; *by = blk / bpr; return; }
  const uint32_t lw = swz & 15u, lh = swz >> 4, per = 1u << (lw + lh);
  const uint32_t st = blk / per, in = blk & (per - 1u), stpr = bpr >> lw;      // bpr is a multiple of 2^lw here
  *bx = ((st % stpr) << lw) + (in & ((1u << lw) - 1u));
  *by = ((st / stpr) << lh) + (in >> lw);
}

// slot base (multiple of 32) -> its tile, the slot index inside the tile
__device__ __forceinline__ int tile_of_slot(const RenderParams& P, uint32_t base, uint32_t* local) {
  int t = 0;
  while (t + 1 < P.ntiles && base >= P.slot_end[t]) ++t;
  *local = base - (t ? P.slot_end[t - 1] : 0u);
  return t;
}

struct Pending { vec3 o, d, w; int depth; };

// CTA shape: A/B on B200 (profiles/r01f_ab_occupancy.txt): 128 threads x 5 CTAs/SM (20 warps, 92-96 registers, no
// spills) beat 256 x 2 (16 warps) by 9 % on C3 and 4 % on C2; 6-8 CTAs/SM (<= 80 / 64 registers) spill and lose.
#ifndef SPRAY_RENDER_MINBLOCKS
#define SPRAY_RENDER_MINBLOCKS 5
#endif
#ifndef SPRAY_RENDER_THREADS
#define SPRAY_RENDER_THREADS 128
#endif
#define SPRAY_RENDER_WARPS (SPRAY_RENDER_THREADS / 32)
static_assert(SPRAY_RENDER_WARPS <= SPRAY_MAX_CTA_WARPS, "CoopScratch (traverse.cuh) is sized for SPRAY_MAX_CTA_WARPS warps per CTA");

// ------------------------------------------------------------------------------------------------
// Per-warp ray queue (shared memory, SoA, LIFO).  Every ray that enters at least one resident domain
// — eye, bounce or shadow — is pushed with its exact domain mask; the warp pops 32 at a time, so all
// lanes traverse (ballot / prefix-sum compaction of the speculative ray stream).  Rays that enter no
// domain never occupy a lane.
// fields (QCAP entries each): 0-2 origin, 3-5 direction, 6-8 throughput (radiance) or radiance to add
// (shadow), 9 pixid, 10 samid, 11 depth << 1 | kind (0 radiance, 1 shadow)
// ------------------------------------------------------------------------------------------------
#define QF_O 0
#define QF_D 3
#define QF_W 6
#define QF_PIX 9
#define QF_SAM 10
#define QF_DK 11
#define QF_MASK 12

struct RayQueue {
  uint32_t* q;   // this warp's fields, field f entry s at q[f * cap + s]
  int cap;
};

__host__ __device__ inline size_t queue_words_per_warp(int cap) { return (size_t)QF_MASK * cap; }

// Warp-cooperative push (ballot / prefix-sum compaction).  A ray that misses every top-level box is not
// queued: the function returns true for those lanes and the caller resolves the ray on the spot
// (radiance: miss, shadow: unoccluded).  `count` is warp-uniform.
__device__ __forceinline__ bool queue_push(const RayQueue& Q, const DomSmem& D, int& count, bool emit, vec3 o, vec3 d,
                                           vec3 w, uint32_t pixid, uint32_t samid, uint32_t depthkind) {
  const int lane = threadIdx.x & 31;
  const bool go = emit && tlas_any(D, make_ray_ctx(o, d, SPRAY_RAY_EPS));
  const uint32_t b = __ballot_sync(0xffffffffu, go);
  if (count + __popc(b) > Q.cap) __trap();   // the queue bound is an invariant of the callers
  if (go) {
    const int s = count + __popc(b & ((1u << lane) - 1u));
    uint32_t* q = Q.q + s;
    const int c = Q.cap;
    q[(QF_O + 0) * c] = __float_as_uint(o.x); q[(QF_O + 1) * c] = __float_as_uint(o.y); q[(QF_O + 2) * c] = __float_as_uint(o.z);
    q[(QF_D + 0) * c] = __float_as_uint(d.x); q[(QF_D + 1) * c] = __float_as_uint(d.y); q[(QF_D + 2) * c] = __float_as_uint(d.z);
    q[(QF_W + 0) * c] = __float_as_uint(w.x); q[(QF_W + 1) * c] = __float_as_uint(w.y); q[(QF_W + 2) * c] = __float_as_uint(w.z);
    q[QF_PIX * c] = pixid; q[QF_SAM * c] = samid; q[QF_DK * c] = depthkind;
  }
  count += __popc(b);
  __syncwarp();
  return emit && !go;
}

// GENERAL == false is the lean variant for the common configuration (PT shader, ONE point light, every
// domain diffuse): area-light / AO sampling, delta materials and the inline shadow fallback are compiled
// out, which keeps the code the instruction cache has to hold small.  The launcher picks the variant.
// Queue bound.  The queue is LIFO, a pop takes 32 rays and eye rays are added only below 32 queued, so
//   lean / no glass : a radiance ray has at most ONE radiance child -> <= 63 radiance entries + 32 shadow rays of the
//                     current pop = 95 (shadow pushes of further light passes fall back to tracing in place);
//   glass           : reflection AND transmission per hit (the reference CHECK-fails there, ooc_shader_pt.h:199-201;
//                     this kernel traces both): every level of the depth-first walk can leave 32 siblings behind
//                     -> 63 + 32 * (bounces - 1) radiance entries + 32 shadow.  The launcher sizes the GENERAL queue
//                     from `bounces` when a glass domain exists (general_queue_cap) and refuses what cannot fit.
template <bool GENERAL> struct QueueCap { static constexpr int value = GENERAL ? 128 : 96; };
static inline int general_queue_cap(bool has_glass, int bounces) {
  if (!has_glass) return QueueCap<true>::value;
  const int need = 63 + 32 * (bounces > 1 ? bounces - 1 : 0) + 32;
  return need < QueueCap<true>::value ? QueueCap<true>::value : (need + 31) / 32 * 32;
}

template <bool COUNT, bool GENERAL>
__global__ void __launch_bounds__(SPRAY_RENDER_THREADS, SPRAY_RENDER_MINBLOCKS) render_tile_kernel(RenderParams P) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(P.table, P.nd, P.tlas, P.tlas_n, smem);
  const int QCAP = GENERAL ? P.qcap : QueueCap<false>::value;   // a compile-time constant in the lean variant
  const int lane = threadIdx.x & 31;
  RayQueue Q;
  Q.cap = QCAP;
  {
    uint32_t* qbase = (uint32_t*)((char*)smem + ((dom_smem_bytes(P.nd, P.tlas_n, 0) + 15) & ~(size_t)15));
    Q.q = qbase + (size_t)(threadIdx.x >> 5) * queue_words_per_warp(QCAP);
  }
  const int spp = P.shade.pixel_samples;
  const float scale = (float)(1.0 / (double)spp);
  const bool pt = GENERAL ? (P.shade.shader == SPRAYGPU_SHADER_PT) : true;
  const int nl = pt ? P.shade.nlights : 1;
  uint32_t cnt[3] = {0, 0, 0};
  uint32_t n_eye = 0, n_rad = 0, n_shadow = 0, n_hit = 0;
  int count = 0;      // rays in this warp's queue (warp-uniform)
  bool more = true;   // the tile still has unclaimed sample slots
  uint32_t claim_next = 0, claim_end = 0;   // this warp's claimed slot range

  for (;;) {
    if (count < 32 && more) {
      // ---------------- refill: one 8x4 pixel block of eye rays (ooc_tracer.inl:84-172) ----------------
      // the warp claims SPRAY_CLAIM_BLOCKS blocks per atomic on the frame's work counter and walks them one at a time
      if (claim_next >= claim_end) {
        unsigned long long b = 0;
        if (lane == 0) b = atomicAdd(P.counters, (unsigned long long)(32 * SPRAY_CLAIM_BLOCKS));
        b = __shfl_sync(0xffffffffu, b, 0);
        if (b >= P.total_slots) { more = false; continue; }
        claim_next = (uint32_t)b;
        claim_end = (uint32_t)(b + 32 * SPRAY_CLAIM_BLOCKS < P.total_slots ? b + 32 * SPRAY_CLAIM_BLOCKS : P.total_slots);
      }
      const unsigned long long base = claim_next;
      claim_next += 32u;
      uint32_t local;
      const int ti = tile_of_slot(P, (uint32_t)base, &local);
      const spraygpu_tile tile = P.tiles[ti];
      const uint32_t bpr = P.blocks_per_row[ti];
      const uint32_t slot = local + lane;
      const uint32_t pix_slot = slot / (uint32_t)spp, sidx = slot % (uint32_t)spp;
      const uint32_t blk = pix_slot >> 5, within = pix_slot & 31u;
      uint32_t bx, by;
      block_xy(blk, bpr, P.swz[ti], &bx, &by);
      const uint32_t x0 = bx * 8u + (within & 7u);
      const uint32_t y0 = by * 4u + (within >> 3);
      const bool has_ray = x0 < (uint32_t)tile.w && y0 < (uint32_t)tile.h;
      uint32_t samid = 0, pixid = 0;
      vec3 dir = mk3(0.f, 0.f, 1.f);
      if (has_ray) {
        eye_sample(P.cam, tile, P.parent_w[ti], P.parent_dx[ti], P.parent_dy[ti], spp, P.shade.eye_seed, x0, y0, sidx, &samid, &pixid, &dir);
        ++n_eye; ++n_rad;
      }
      queue_push(Q, D, count, has_ray, mk3(P.cam.pos[0], P.cam.pos[1], P.cam.pos[2]), dir, mk3(1.f, 1.f, 1.f), pixid,
                 samid, 0u);
      continue;
    }
    if (count == 0) break;

    // ---------------- pop up to 32 rays; every popped lane traverses ----------------
    const int n = count < 32 ? count : 32;
    count -= n;
    const bool act = lane < n;
    const uint32_t* qe = Q.q + count + (act ? lane : 0);
    const vec3 ro = mk3(__uint_as_float(qe[(QF_O + 0) * QCAP]), __uint_as_float(qe[(QF_O + 1) * QCAP]),
                        __uint_as_float(qe[(QF_O + 2) * QCAP]));
    const vec3 rd = mk3(__uint_as_float(qe[(QF_D + 0) * QCAP]), __uint_as_float(qe[(QF_D + 1) * QCAP]),
                        __uint_as_float(qe[(QF_D + 2) * QCAP]));
    const uint32_t dk = qe[QF_DK * QCAP];
    const bool shadow = (dk & 1u) != 0u;
    SceneHit h; h.t = INFINITY; h.dom = 0; h.prim = 0; h.u = h.v = 0.f;
    bool found = false;
    found = scene_trace_warp<COUNT>(P.table, D, act, ro, rd, shadow, h, cnt);
    // payload (read before any push overwrites the popped entries)
    const vec3 Lin = mk3(__uint_as_float(qe[(QF_W + 0) * QCAP]), __uint_as_float(qe[(QF_W + 1) * QCAP]),
                         __uint_as_float(qe[(QF_W + 2) * QCAP]));
    const uint32_t pixid = qe[QF_PIX * QCAP], samid = qe[QF_SAM * QCAP];
    __syncwarp();

    if (act && shadow && !found) fb_add(P.fb, pixid, scale, Lin);

    // ---------------- shade radiance hits (ooc_shader_pt.h:94-227, ooc_shader_ao.h:92-197) ----------------
    const bool hit = act && !shadow && found;
    const int next_depth = (int)(dk >> 1) + 1;
    HitCtx c;
    c.pos = ro; c.kd = mk3(0, 0, 0); c.nff = mk3(0, 0, 1); c.wo = gneg(rd);
    c.cos_theta_i = 0.f; c.entering = false; c.mtl = SPRAYGPU_MTL_DIFFUSE; c.eta0 = c.eta1 = 1.f;
    if (hit) {
      ++n_hit;
      make_hit_ctx<GENERAL>(P.table + h.dom, ro, rd, h.t, h.u, h.v, h.prim, c);
    }

    // ---------------- secondary rays: pushed first (the queue has room for them by construction: QueueCap) ----------------
    {
      bool em1 = false, em2 = false;
      vec3 d1 = mk3(0, 0, 1), w1 = mk3(0, 0, 0), d2 = mk3(0, 0, 1), w2 = mk3(0, 0, 0);
      if (hit && next_depth < P.shade.bounces) shade_secondary<GENERAL>(c, Lin, samid, next_depth, &em1, &d1, &w1, &em2, &d2, &w2);
      if (em1) ++n_rad;
      if (__any_sync(0xffffffffu, em1))
        queue_push(Q, D, count, em1, c.pos, d1, w1, pixid, samid, (uint32_t)next_depth << 1);
      if (GENERAL) {
        if (em2) ++n_rad;
        if (__any_sync(0xffffffffu, em2))
          queue_push(Q, D, count, em2, c.pos, d2, w2, pixid, samid, (uint32_t)next_depth << 1);
      }
    }
    // ---------------- shadow rays: one light sample per lane per pass ----------------
    {
      uint32_t ls = rng_init1(samid * (uint32_t)next_depth);  // ShaderPt light_sampler (ooc_shader_pt.h:129-130)
      int l = 0, k = 0;
      bool more_lights = hit && c.mtl == SPRAYGPU_MTL_DIFFUSE && nl > 0;
      while (__any_sync(0xffffffffu, more_lights)) {
        bool emit = false; vec3 wi = mk3(0, 0, 1), Lr = mk3(0, 0, 0);
        if (more_lights) emit = shade_light_pass<GENERAL>(P.shade, pt, nl, c, Lin, pixid, l, k, ls, &wi, &Lr, &more_lights);
        if (emit) ++n_shadow;
        if (!GENERAL || count + 32 <= QCAP) {
          // a shadow ray that enters no resident domain is unoccluded
          if (queue_push(Q, D, count, emit, c.pos, wi, Lr, pixid, samid, 1u)) fb_add(P.fb, pixid, scale, Lr);
        } else {
          // queue full (many light samples per hit): trace this pass in place
          SceneHit hs;
          const bool oc = scene_trace_warp<COUNT>(P.table, D, emit, c.pos, wi, true, hs, cnt);
          if (emit && !oc) fb_add(P.fb, pixid, scale, Lr);
        }
      }
    }
  }

  // statistics: warp-reduce, one atomic per warp per counter
  uint32_t vals[7] = {n_eye, n_rad, n_shadow, cnt[0], cnt[1], cnt[2], n_hit};
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    unsigned long long v = vals[k];
#pragma unroll
    for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0 && v) atomicAdd(P.counters + 1 + k, v);
  }
}

// ------------------------------------------------------------------------------------------------
// The STREAMING variant of the fused kernel (lean configuration: PT shader, one point light, diffuse
// domains — the reference's example scenes).
//
// render_tile_kernel above traces rays in batches of 32: every lane waits for the slowest ray of its batch
// (ncu r01e: 8 of 32 lanes active in the node test on incoherent bounce rays).  Here a lane whose ray is
// finished does not wait: as soon as SPRAY_FETCH_THRESH lanes are idle the warp runs ONE service pass —
// the idle lanes shade their finished rays, push the bounce and shadow rays (ballot / prefix-sum
// compaction into the warp's shared-memory queue), the queue is topped up with eye rays, and the idle
// lanes pop new rays — while the other lanes keep their traversal state in registers.  The traversal
// loop itself is the same LaneTrav state machine (one node step + one triangle step per pass).
//
// Queue bound (QCAP = 128): a ray lineage owns at most one radiance ray; shadow rays are pushed last and
// at least as many rays are popped right after (n_idle >= shadows pushed), so shadow rays never pile up;
// eye rays are only added when fewer than 32 rays are queued.  => queued <= 95 radiance + 32 transient.
// ------------------------------------------------------------------------------------------------
#ifndef SPRAY_FETCH_THRESH
#define SPRAY_FETCH_THRESH 12
#endif
#define STREAM_QCAP 128
// per-lane in-flight record (shared memory, field f of lane l at base[f * 32 + l]): the popped ray's payload
#define IF_W 0
#define IF_PIX 3
#define IF_SAM 4
#define IF_DK 5
#define IF_WORDS 6

__host__ __device__ inline size_t stream_words_per_warp() { return queue_words_per_warp(STREAM_QCAP) + (size_t)IF_WORDS * 32; }

template <bool COUNT>
__global__ void __launch_bounds__(SPRAY_RENDER_THREADS, SPRAY_RENDER_MINBLOCKS) render_tile_stream_kernel(RenderParams P) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(P.table, P.nd, P.tlas, P.tlas_n, smem);
  const int lane = threadIdx.x & 31;
  RayQueue Q;
  Q.cap = STREAM_QCAP;
  uint32_t* inflight;
  {
    uint32_t* qbase = (uint32_t*)((char*)smem + ((dom_smem_bytes(P.nd, P.tlas_n, 0) + 15) & ~(size_t)15));
    Q.q = qbase + (size_t)(threadIdx.x >> 5) * stream_words_per_warp();
    inflight = Q.q + queue_words_per_warp(STREAM_QCAP) + lane;
  }
  const int spp = P.shade.pixel_samples;
  const float scale = (float)(1.0 / (double)spp);
  uint32_t cnt[3] = {0, 0, 0};
  uint32_t n_eye = 0, n_rad = 0, n_shadow = 0, n_hit = 0;
  int count = 0;      // rays in this warp's queue (warp-uniform)
  bool more = true;   // the tile still has unclaimed sample slots (warp-uniform)

  LaneStack stack;
  LaneTrav T(stack);
  T.begin(false);
  TlasStack tstack;
  TlasSource src(P.table, D, tstack);
  RayCtx r = make_ray_ctx(mk3(0, 0, 0), mk3(0, 0, 1), SPRAY_RAY_EPS);
  bool busy = false;  // this lane holds a ray (being traced, or finished and not yet shaded)
  bool any = false;   // the held ray is a shadow ray

  for (;;) {
    __syncwarp();
    // lane state: running (has a node or a triangle to test), waiting (needs its next domain from the top-level
    // BVH: done in the service pass, all waiting lanes together), finished (ray done, to be shaded), empty
#if SPRAY_POP_CULL
    T.oct = r.octinv;
#endif
    int w = T.want();
    const bool running = w == LaneTrav::WANT_NODE || w == LaneTrav::WANT_TRI;
    const bool waiting = w == LaneTrav::WANT_SOURCE;
    const bool fin = busy && !T.alive;
    const uint32_t rm = __ballot_sync(0xffffffffu, running);
    const bool pending = __any_sync(0xffffffffu, waiting || fin) || count > 0 || more;
    if (rm == 0u && !pending) break;   // nothing in flight, nothing queued, tile exhausted
    if (pending && (rm == 0u || 32 - __popc(rm) >= SPRAY_FETCH_THRESH)) {
      // ================= service pass =================
      const vec3 Lin = mk3(__uint_as_float(inflight[(IF_W + 0) * 32]), __uint_as_float(inflight[(IF_W + 1) * 32]),
                           __uint_as_float(inflight[(IF_W + 2) * 32]));
      const uint32_t pixid = inflight[IF_PIX * 32], samid = inflight[IF_SAM * 32], dk = inflight[IF_DK * 32];
      // ---- A. retire shadow rays, shade radiance hits (ooc_shader_pt.h:94-227)
      if (fin && any && !T.found) fb_add(P.fb, pixid, scale, Lin);
      const bool hit = fin && !any && T.out.t < INFINITY;
      const int next_depth = (int)(dk >> 1) + 1;
      HitCtx c;
      c.pos = r.o; c.kd = mk3(0, 0, 0); c.nff = mk3(0, 0, 1); c.wo = gneg(r.d);
      c.cos_theta_i = 0.f; c.entering = false; c.mtl = SPRAYGPU_MTL_DIFFUSE; c.eta0 = c.eta1 = 1.f;
      bool em1 = false, em2 = false, emit = false;
      vec3 d1 = mk3(0, 0, 1), w1 = mk3(0, 0, 0), d2 = mk3(0, 0, 1), w2 = mk3(0, 0, 0), wi = mk3(0, 0, 1), Lr = mk3(0, 0, 0);
      if (hit) {
        ++n_hit;
        make_hit_ctx<false>(P.table + T.out.dom, r.o, r.d, T.out.t, T.out.u, T.out.v, T.out.prim, c);
        if (next_depth < P.shade.bounces) shade_secondary<false>(c, Lin, samid, next_depth, &em1, &d1, &w1, &em2, &d2, &w2);
        uint32_t ls = rng_init1(samid * (uint32_t)next_depth);  // ShaderPt light_sampler (ooc_shader_pt.h:129-130)
        int l = 0, k = 0; bool more_lights = true;
        emit = shade_light_pass<false>(P.shade, true, 1, c, Lin, pixid, l, k, ls, &wi, &Lr, &more_lights);
      }
      if (fin) busy = false;
      // ---- B. push: bounce rays first, shadow rays on top (popped first)
      if (em1) ++n_rad;
      if (__any_sync(0xffffffffu, em1))
        queue_push(Q, D, count, em1, c.pos, d1, w1, pixid, samid, (uint32_t)next_depth << 1);
      if (emit) ++n_shadow;
      if (__any_sync(0xffffffffu, emit)) {
        // a shadow ray that enters no resident domain is unoccluded
        if (queue_push(Q, D, count, emit, c.pos, wi, Lr, pixid, samid, 1u)) fb_add(P.fb, pixid, scale, Lr);
      }
      // ---- C. top up with one 8x4 pixel block of eye rays (ooc_tracer.inl:84-172)
      if (count < 32 && more) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(P.counters, 32ull);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= P.total_slots) more = false;
        else {
          uint32_t local;
          const int ti = tile_of_slot(P, (uint32_t)base, &local);
          const spraygpu_tile tile = P.tiles[ti];
          const uint32_t bpr = P.blocks_per_row[ti];
          const uint32_t slot = local + lane;
          const uint32_t pix_slot = slot / (uint32_t)spp, sidx = slot % (uint32_t)spp;
          const uint32_t blk = pix_slot >> 5, within = pix_slot & 31u;
          uint32_t bx, by;
          block_xy(blk, bpr, P.swz[ti], &bx, &by);
          const uint32_t x0 = bx * 8u + (within & 7u);
          const uint32_t y0 = by * 4u + (within >> 3);
          const bool has_ray = x0 < (uint32_t)tile.w && y0 < (uint32_t)tile.h;
          uint32_t esam = 0, epix = 0;
          vec3 dir = mk3(0.f, 0.f, 1.f);
          if (has_ray) {
            eye_sample(P.cam, tile, P.parent_w[ti], P.parent_dx[ti], P.parent_dy[ti], spp, P.shade.eye_seed, x0, y0, sidx, &esam, &epix, &dir);
            ++n_eye; ++n_rad;
          }
          queue_push(Q, D, count, has_ray, mk3(P.cam.pos[0], P.cam.pos[1], P.cam.pos[2]), dir, mk3(1.f, 1.f, 1.f), epix,
                     esam, 0u);
        }
      }
      // ---- D. empty lanes pop new rays (top of the queue first)
      {
        const bool idle = !busy;
        const uint32_t im = __ballot_sync(0xffffffffu, idle);
        const int nf = min(__popc(im), count);
        const int rnk = __popc(im & ((1u << lane) - 1u));
        const bool take = idle && rnk < nf;
        if (take) {
          const uint32_t* qe = Q.q + (count - 1 - rnk);
          const vec3 ro = mk3(__uint_as_float(qe[(QF_O + 0) * STREAM_QCAP]), __uint_as_float(qe[(QF_O + 1) * STREAM_QCAP]),
                              __uint_as_float(qe[(QF_O + 2) * STREAM_QCAP]));
          const vec3 rd = mk3(__uint_as_float(qe[(QF_D + 0) * STREAM_QCAP]), __uint_as_float(qe[(QF_D + 1) * STREAM_QCAP]),
                              __uint_as_float(qe[(QF_D + 2) * STREAM_QCAP]));
#pragma unroll
          for (int f = 0; f < 6; ++f) inflight[(IF_W + f) * 32] = qe[(QF_W + f) * STREAM_QCAP];
          any = (qe[QF_DK * STREAM_QCAP] & 1u) != 0u;
          r = make_ray_ctx(ro, rd, SPRAY_RAY_EPS);
          busy = true;
          src.reset(true);             // the new ray starts at the top-level root (other lanes keep their walk)
          T.begin(true);
        }
        count -= nf;
      }
      __syncwarp();
      // ---- E. top-level BVH steps for every lane that needs its next domain (old and new), together
#if SPRAY_POP_CULL
      T.oct = r.octinv;   // r may have just been replaced by a new ray
#endif
      w = T.want();
      if (w == LaneTrav::WANT_SOURCE) { T.source_step<COUNT>(src, r, cnt); w = T.want(); }
      __syncwarp();
    }
    // ================= traversal pass: one node step + one triangle step per lane =================
    if (w == LaneTrav::WANT_NODE) T.node_step<COUNT>(r, cnt);
    __syncwarp();
#pragma unroll 1
    for (int k = 0; k < SPRAY_TRI_PER_PASS && T.alive && T.tgy; ++k) T.tri_step<COUNT>(r, any, cnt);
  }

  uint32_t vals[7] = {n_eye, n_rad, n_shadow, cnt[0], cnt[1], cnt[2], n_hit};
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    unsigned long long v = vals[k];
#pragma unroll
    for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0 && v) atomicAdd(P.counters + 1 + k, v);
  }
}

// ------------------------------------------------------------------------------------------------
// host-side launchers (called from capi.cu)
// ------------------------------------------------------------------------------------------------
static inline int grid_for(size_t n, int t) { return (int)((n + t - 1) / t); }

CameraDev to_camdev(const spraygpu_camera* c) {
  CameraDev d;
  for (int k = 0; k < 3; ++k) { d.pos[k] = c->pos[k]; d.lowerleft[k] = c->lowerleft[k]; d.wvec[k] = c->wvec[k]; d.hvec[k] = c->hvec[k]; }
  d.image_w = c->image_w; d.image_h = c->image_h;
  return d;
}

int launch_domain_lists(Device* dev, const spraygpu_ray* rays, size_t n, int cap, int32_t* counts, int32_t* ids,
                        float* tmins) {
  if (!n) return SPRAYGPU_OK;
  SG_DYN_SMEM(domain_lists_kernel, dom_smem_bytes(dev->ndomains, dev->tlas_n, 4));
  domain_lists_kernel<<<grid_for(n, 128), 128, dom_smem_bytes(dev->ndomains, dev->tlas_n, 4), dev->stream>>>(
      dev->d_domains, dev->ndomains, dev->d_tlas, dev->tlas_n, rays, n, cap, counts, ids, tmins);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}
int launch_intersect_scene(Scene* sc, int domain_id, const spraygpu_ray* rays, size_t n, uint64_t* keys, float* uv) {
  if (!n) return SPRAYGPU_OK;
  intersect_scene_kernel<<<grid_for(n, 128), 128, 0, sc->dev->stream>>>(sc->d_nodes, sc->d_tris, (uint32_t)domain_id,
                                                                        rays, n, keys, uv);
  sc->dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}
int launch_occluded_scene(Scene* sc, const spraygpu_ray* rays, size_t n, uint8_t* occ) {
  if (!n) return SPRAYGPU_OK;
  occluded_scene_kernel<<<grid_for(n, 128), 128, 0, sc->dev->stream>>>(sc->d_nodes, sc->d_tris, rays, n, occ);
  sc->dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}
int launch_intersect_domains(Device* dev, const spraygpu_ray* rays, size_t n, uint64_t* keys, float* uv, int merge) {
  if (!n) return SPRAYGPU_OK;
  SG_DYN_SMEM(intersect_domains_kernel, dom_smem_bytes(dev->ndomains, dev->tlas_n, 0));
  intersect_domains_kernel<<<grid_for(n, 128), 128, dom_smem_bytes(dev->ndomains, dev->tlas_n, 0), dev->stream>>>(
      dev->d_domains, dev->ndomains, dev->d_tlas, dev->tlas_n, rays, n, keys, uv, merge);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}
int launch_occluded_domains(Device* dev, const spraygpu_ray* rays, size_t n, uint8_t* occ, int merge) {
  if (!n) return SPRAYGPU_OK;
  SG_DYN_SMEM(occluded_domains_kernel, dom_smem_bytes(dev->ndomains, dev->tlas_n, 0));
  occluded_domains_kernel<<<grid_for(n, 128), 128, dom_smem_bytes(dev->ndomains, dev->tlas_n, 0), dev->stream>>>(
      dev->d_domains, dev->ndomains, dev->d_tlas, dev->tlas_n, rays, n, occ, merge);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}
int launch_hit_attributes(Device* dev, const spraygpu_ray* rays, const uint64_t* keys, size_t n, uint32_t* color,
                          float* ns3) {
  if (!n) return SPRAYGPU_OK;
  hit_attributes_kernel<<<grid_for(n, 128), 128, 0, dev->stream>>>(dev->d_domains, rays, keys, n, color, ns3);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}
int launch_eye_rays(Device* dev, const spraygpu_camera* cam, spraygpu_tile tile, int spp, int eye_seed, spraygpu_ray* rays) {
  const size_t n = (size_t)tile.w * tile.h * spp;
  if (!n) return SPRAYGPU_OK;
  eye_rays_kernel<<<grid_for(n, 256), 256, 0, dev->stream>>>(to_camdev(cam), tile, spp, eye_seed, rays);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

int launch_render_tile(Device* dev, const spraygpu_camera* cam, const spraygpu_shade* shade, spraygpu_tile tile,
                       float* fb, bool counted) {
  return launch_render_tiles(dev, cam, shade, &tile, nullptr, 1, fb, counted, true);
}

// up to SPRAY_MAX_FRAME_TILES tiles in ONE persistent launch; reset_counters = false keeps accumulating the
// statistics of a frame that needs several launches
// parents == nullptr: every tile is a whole blocking tile; else tiles[t] is a sub-rectangle of the blocking tile parents[t]
// SPRAY_BLOCK_SWIZZLE=k (A/B knob, read once): enumerate a tile's pixel blocks in super-tiles of 2^k x 2^(k+1) blocks
static int swizzle_log2() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SPRAY_BLOCK_SWIZZLE"); v = e ? atoi(e) : SPRAY_BLOCK_SWIZZLE_DEFAULT; if (v < 0 || v > 7) v = 0; }
  return v;
}

int launch_render_tiles(Device* dev, const spraygpu_camera* cam, const spraygpu_shade* shade, const spraygpu_tile* tiles,
                        const spraygpu_tile* parents, int ntiles, float* fb, bool counted, bool reset_counters) {
  SG_CHECK(ntiles >= 0 && ntiles <= SPRAY_MAX_FRAME_TILES, SPRAYGPU_ERR_LIMIT, "%d tiles in one launch (max %d)", ntiles, SPRAY_MAX_FRAME_TILES);
  RenderParams P;
  P.table = dev->d_domains; P.nd = dev->ndomains; P.tlas = dev->d_tlas; P.tlas_n = dev->tlas_n;
  P.cam = to_camdev(cam); P.shade = *shade; P.fb = fb;
  P.counters = dev->d_counters;
  uint64_t slots = 0;
  P.ntiles = 0;
  for (int t = 0; t < ntiles; ++t) {
    const spraygpu_tile tile = tiles[t];
    uint32_t bpr = (uint32_t)((tile.w + 7) / 8), brows = (uint32_t)((tile.h + 3) / 4);
    uint8_t swz = 0;
    if (swizzle_log2() > 0 && bpr && brows) {
      // super-tiles of up to 2^k x 2^(k+1) blocks (8 x 4 pixels each: 64 x 64 pixels at k = 3), shrunk to the tile
      uint32_t lw = (uint32_t)swizzle_log2(), lh = lw + 1;
      while (lw > 0 && (1u << lw) > bpr) --lw;
      while (lh > 0 && (1u << lh) > brows) --lh;
      swz = (uint8_t)(lw | (lh << 4));
      bpr = ((bpr + (1u << lw) - 1) >> lw) << lw;
      brows = ((brows + (1u << lh) - 1) >> lh) << lh;
    }
    const uint64_t ts = (uint64_t)bpr * (uint64_t)brows * 32ull * (uint64_t)shade->pixel_samples;
    if (!ts) continue;
    slots += ts;
    SG_CHECK(slots < 0xFFFFFFFFull, SPRAYGPU_ERR_LIMIT, "tiles too large: %llu sample slots", (unsigned long long)slots);
    P.tiles[P.ntiles] = tile; P.blocks_per_row[P.ntiles] = bpr; P.slot_end[P.ntiles] = (uint32_t)slots; P.swz[P.ntiles] = swz;
    P.parent_w[P.ntiles] = tile.w; P.parent_dx[P.ntiles] = 0; P.parent_dy[P.ntiles] = 0;
    if (parents) {
      const spraygpu_tile par = parents[t];
      SG_CHECK(tile.x >= par.x && tile.y >= par.y && tile.x + tile.w <= par.x + par.w && tile.y + tile.h <= par.y + par.h &&
                   tile.x - par.x < 65536 && tile.y - par.y < 65536,
               SPRAYGPU_ERR_ARG, "rectangle %d (%d,%d,%d,%d) is not inside its blocking tile (%d,%d,%d,%d)", t, tile.x, tile.y, tile.w,
               tile.h, par.x, par.y, par.w, par.h);
      P.parent_w[P.ntiles] = par.w; P.parent_dx[P.ntiles] = (uint16_t)(tile.x - par.x); P.parent_dy[P.ntiles] = (uint16_t)(tile.y - par.y);
    }
    ++P.ntiles;
  }
  for (int t = P.ntiles; t < SPRAY_MAX_FRAME_TILES; ++t) {
    P.tiles[t] = spraygpu_tile{0, 0, 0, 0}; P.blocks_per_row[t] = 1; P.slot_end[t] = (uint32_t)slots; P.swz[t] = 0;
    P.parent_w[t] = 0; P.parent_dx[t] = 0; P.parent_dy[t] = 0;
  }
  P.total_slots = (uint32_t)slots;
  if (reset_counters) SG_CUDA(cudaMemsetAsync(dev->d_counters, 0, sizeof(unsigned long long) * 8, dev->stream));
  else SG_CUDA(cudaMemsetAsync(dev->d_counters, 0, sizeof(unsigned long long), dev->stream));   // the work counter only
  if (!slots) return SPRAYGPU_OK;
  // lean variant: PT shader, ONE point light, all domains diffuse
  bool general = shade->shader != SPRAYGPU_SHADER_PT || shade->nlights != 1;
  for (int i = 0; i < shade->nlights; ++i) general = general || shade->lights[i].type != SPRAYGPU_LIGHT_POINT;
  for (int i = 0; i < dev->ndomains; ++i) general = general || dev->h_domains[i].mtl_type != SPRAYGPU_MTL_DIFFUSE;
  // The streaming variant is opt-in (SPRAYGPU_STREAM=1): on B200 it lost to the batch-of-32 kernel on both
  // measured workloads (C2 2.7-3.2 vs 4.8 Grays/s, C3 0.76-0.81 vs 0.85 Grays/s).
  static const bool want_stream = getenv("SPRAYGPU_STREAM") && atoi(getenv("SPRAYGPU_STREAM")) != 0;
  const bool stream = !general && want_stream;
  void (*kern)(RenderParams) =
      stream ? (counted ? render_tile_stream_kernel<true> : render_tile_stream_kernel<false>)
             : counted ? (general ? render_tile_kernel<true, true> : render_tile_kernel<true, false>)
                       : (general ? render_tile_kernel<false, true> : render_tile_kernel<false, false>);
  bool has_glass = false;
  for (int i = 0; i < dev->ndomains; ++i) has_glass = has_glass || dev->h_domains[i].mtl_type == SPRAYGPU_MTL_GLASS;
  const int qcap = general ? general_queue_cap(has_glass, shade->bounces) : QueueCap<false>::value;
  P.qcap = qcap;
  const size_t smem = ((dom_smem_bytes(dev->ndomains, dev->tlas_n, 0) + 15) & ~(size_t)15) +
                      SPRAY_RENDER_WARPS * (stream ? stream_words_per_warp() : queue_words_per_warp(qcap)) * sizeof(uint32_t);
  // attribute + occupancy are per (kernel, shared-memory size): cached, the frame loop is host-latency sensitive at
  // 8 GPUs (0.1 ms of kernel per rank)
  int per_sm = dev->render_per_sm;
  if (dev->render_kern != (const void*)kern || dev->render_smem != smem) {
    SG_CHECK(smem <= 227 * 1024, SPRAYGPU_ERR_LIMIT, "render: %zu B of shared memory per CTA (glass domains with %d bounces need a "
             "%d-entry ray queue per warp); reduce --bounces", smem, shade->bounces, qcap);
    SG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, SPRAY_RENDER_THREADS, smem));
    if (per_sm < 1) per_sm = 1;
    dev->render_kern = (const void*)kern; dev->render_smem = smem; dev->render_per_sm = per_sm;
  }
  // persistent grid: SM count x resident CTAs per SM
  int grid = dev->sm_count * per_sm;
  const uint64_t warps_needed = (slots + 31) / 32;
  if ((uint64_t)grid * SPRAY_RENDER_WARPS > warps_needed) grid = (int)((warps_needed + SPRAY_RENDER_WARPS - 1) / SPRAY_RENDER_WARPS);
  kern<<<grid, SPRAY_RENDER_THREADS, smem, dev->stream>>>(P);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

}  // namespace spraygpu
