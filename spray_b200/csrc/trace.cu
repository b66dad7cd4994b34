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
#include "internal.cuh"

namespace spraygpu {

#define MAXLIST 16
#define MAX_PENDING 8

struct NodeLoaderG {
  const uint4* nodes;
  __device__ __forceinline__ void operator()(uint32_t idx, u4& q0, u4& q1, u4& q2, u4& q3, u4& q4) const {
    const uint4* p = nodes + 5 * (size_t)idx;
    const uint4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3), e = __ldg(p + 4);
    q0.x = a.x; q0.y = a.y; q0.z = a.z; q0.w = a.w;
    q1.x = b.x; q1.y = b.y; q1.z = b.z; q1.w = b.w;
    q2.x = c.x; q2.y = c.y; q2.z = c.z; q2.w = c.w;
    q3.x = d.x; q3.y = d.y; q3.z = d.z; q3.w = d.w;
    q4.x = e.x; q4.y = e.y; q4.z = e.z; q4.w = e.w;
  }
};
struct TriLoaderG {
  const float4* tris;
  __device__ __forceinline__ void operator()(uint32_t idx, vec3& v0, vec3& e1, vec3& e2, uint32_t& prim) const {
    const float4* p = tris + 3 * (size_t)idx;
    const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    v0 = mk3(a.x, a.y, a.z); prim = __float_as_uint(a.w);
    e1 = mk3(b.x, b.y, b.z);
    e2 = mk3(c.x, c.y, c.z);
  }
};

// ------------------------------------------------------------------------------------------------
// Domain culling (a6), two stages:
//  1. warp_cull: the warp's active rays are bounded by an origin box and a direction box; every lane tests
//     nd/32 domain boxes against that BEAM (decoupled per-axis interval test, conservative) and a ballot
//     publishes the candidate bit mask.  For coherent rays (8x4 pixel blocks, shadow rays to one light)
//     this leaves ~10 of 64 boxes; for incoherent rays it degrades gracefully to "all boxes".
//  2. each lane runs the reference's exact slab test (aabb.h:139-169) on the candidates only and keeps
//     its own (tmin, id)-sorted list.
// ------------------------------------------------------------------------------------------------
#define CAND_WORDS (SPRAYGPU_MAX_DOMAINS / 32)

struct DomSmem {
  const float* boxes;       // nd x 6
  const uint8_t* resident;  // nd
  uint32_t* cand;           // this warp's CAND_WORDS candidate words
  int nd;
};

__host__ __device__ inline size_t dom_smem_bytes(int nd, int nwarps) {
  return (size_t)nd * 6 * sizeof(float) + (((size_t)nd + 15) & ~(size_t)15) + (size_t)nwarps * CAND_WORDS * 4;
}

__device__ __forceinline__ DomSmem load_domain_smem(const DomainEntry* __restrict__ table, int nd, float* smem) {
  float* boxes = smem;
  uint8_t* resident = (uint8_t*)(boxes + 6 * nd);
  uint32_t* cand = (uint32_t*)(resident + ((nd + 15) & ~15));
  for (int i = threadIdx.x; i < nd; i += blockDim.x) {
#pragma unroll
    for (int k = 0; k < 6; ++k) boxes[6 * i + k] = table[i].box[k];
    resident[i] = (uint8_t)table[i].resident;
  }
  __syncthreads();
  DomSmem d;
  d.boxes = boxes; d.resident = resident; d.cand = cand + (threadIdx.x >> 5) * CAND_WORDS; d.nd = nd;
  return d;
}

__device__ __forceinline__ float warp_min_f(float v) { return sortable_float(__reduce_min_sync(0xffffffffu, float_sortable(v))); }
__device__ __forceinline__ float warp_max_f(float v) { return sortable_float(__reduce_max_sync(0xffffffffu, float_sortable(v))); }

// one axis of the beam/box test: the beam's coordinate range at parameter t >= 0 is
// [o0 + t*d0, o1 + t*d1]; it must overlap [lo, hi]:  o0 + t*d0 <= hi  and  o1 + t*d1 >= lo.
__device__ __forceinline__ void beam_axis(float lo, float hi, float o0, float o1, float d0, float d1, float r0, float r1,
                                          float& tl, float& th) {
  const float a = (hi - o0) * r0;  // o0 + t*d0 <= hi
  if (d0 > 0.0f) th = fminf(th, a);
  else if (d0 < 0.0f) tl = fmaxf(tl, a);
  else if (o0 > hi) th = -INFINITY;
  const float b = (lo - o1) * r1;  // o1 + t*d1 >= lo
  if (d1 > 0.0f) tl = fmaxf(tl, b);
  else if (d1 < 0.0f) th = fminf(th, b);
  else if (o1 < lo) th = -INFINITY;
}

// `only_resident`: skip domains that are not resident on this GPU (traversal); the parity surface
// spraygpu_domain_lists reports all entered domains.
__device__ __forceinline__ void warp_cull(const DomSmem& D, bool active, vec3 o, vec3 d, bool only_resident) {
  const int lane = threadIdx.x & 31;
  const int nwords = (D.nd + 31) >> 5;
  const float o0x = warp_min_f(active ? o.x : INFINITY), o1x = warp_max_f(active ? o.x : -INFINITY);
  const float o0y = warp_min_f(active ? o.y : INFINITY), o1y = warp_max_f(active ? o.y : -INFINITY);
  const float o0z = warp_min_f(active ? o.z : INFINITY), o1z = warp_max_f(active ? o.z : -INFINITY);
  const float d0x = warp_min_f(active ? d.x : INFINITY), d1x = warp_max_f(active ? d.x : -INFINITY);
  const float d0y = warp_min_f(active ? d.y : INFINITY), d1y = warp_max_f(active ? d.y : -INFINITY);
  const float d0z = warp_min_f(active ? d.z : INFINITY), d1z = warp_max_f(active ? d.z : -INFINITY);
  const bool any = o0x <= o1x;  // false when no lane is active (or NaN origins: then pass everything below)
  const bool finite = (o1x - o0x) < INFINITY && (o1y - o0y) < INFINITY && (o1z - o0z) < INFINITY &&
                      (d1x - d0x) < INFINITY && (d1y - d0y) < INFINITY && (d1z - d0z) < INFINITY;
  // scene-scale pad: absorbs the FP32 rounding differences between this test and the exact slab test
  const float pad = 1e-4f * (fabsf(o0x) + fabsf(o1x) + fabsf(o0y) + fabsf(o1y) + fabsf(o0z) + fabsf(o1z) + 1.0f);
  const float r0x = __frcp_rn(d0x), r1x = __frcp_rn(d1x), r0y = __frcp_rn(d0y), r1y = __frcp_rn(d1y),
              r0z = __frcp_rn(d0z), r1z = __frcp_rn(d1z);
  for (int w = 0; w < nwords; ++w) {
    const int i = (w << 5) + lane;
    bool pass = false;
    if (i < D.nd && (!only_resident || D.resident[i])) {
      if (!finite) pass = any;
      else if (any) {
        const float* b = D.boxes + 6 * i;
        const float bp = pad + 1e-4f * (fabsf(b[0]) + fabsf(b[1]) + fabsf(b[2]) + fabsf(b[3]) + fabsf(b[4]) + fabsf(b[5]));
        float tl = 0.0f, th = INFINITY;
        beam_axis(b[0] - bp, b[3] + bp, o0x, o1x, d0x, d1x, r0x, r1x, tl, th);
        beam_axis(b[1] - bp, b[4] + bp, o0y, o1y, d0y, d1y, r0y, r1y, tl, th);
        beam_axis(b[2] - bp, b[5] + bp, o0z, o1z, d0z, d1z, r0z, r1z, tl, th);
        pass = !(tl * 0.9999f > th * 1.0001f + 1e-6f);  // NaN -> pass (conservative)
      }
    }
    const uint32_t m = __ballot_sync(0xffffffffu, pass);
    if (lane == 0) D.cand[w] = m;
  }
  __syncwarp();
}

// candidate-word sources for the per-lane exact test
struct WarpCand {      // the warp's beam-culled candidate words (warp_cull)
  const uint32_t* cand;
  __device__ __forceinline__ uint32_t operator()(int w) const { return cand[w]; }
};
struct StridedCand {   // one ray's own exact domain mask, word w at base[w * stride] (ray queue, SoA)
  const uint32_t* base; int stride;
  __device__ __forceinline__ uint32_t operator()(int w) const { return base[w * stride]; }
};

// per-lane exact test on the candidates: keeps the MAXLIST nearest entries with key > after.
template <typename CandF>
__device__ __forceinline__ int domain_list(const DomSmem& D, const CandF& cand, vec3 o, vec3 d, bool have_after,
                                           uint64_t after, uint64_t* list, bool* more) {
  int cnt = 0;
  bool overflow = false;
  const int nwords = (D.nd + 31) >> 5;
  for (int w = 0; w < nwords; ++w) {
    uint32_t m = cand(w);
    while (m) {
      const int i = (w << 5) + __ffs((int)m) - 1;
      m &= m - 1;
      float tmin;
      if (!domain_slab(D.boxes + 6 * i, o, d, SPRAY_RAY_EPS, INFINITY, &tmin)) continue;
      const uint64_t key = ((uint64_t)float_sortable(tmin) << 32) | (uint32_t)i;
      if (have_after && key <= after) continue;
      if (cnt == MAXLIST) {
        overflow = true;
        if (key >= list[MAXLIST - 1]) continue;
        cnt = MAXLIST - 1;
      }
      int k = cnt;
      while (k > 0 && list[k - 1] > key) { list[k] = list[k - 1]; --k; }
      list[k] = key;
      ++cnt;
    }
  }
  *more = overflow;
  return cnt;
}

// exact per-lane domain mask of candidate word w (reference slab test, aabb.h:139-169)
__device__ __forceinline__ uint32_t exact_mask_word(const DomSmem& D, int w, uint32_t m, vec3 o, vec3 d) {
  uint32_t out = 0;
  while (m) {
    const int b = __ffs((int)m) - 1;
    m &= m - 1;
    float tmin;
    if (domain_slab(D.boxes + 6 * ((w << 5) + b), o, d, SPRAY_RAY_EPS, INFINITY, &tmin)) out |= 1u << b;
  }
  return out;
}

struct SceneHit {
  float t, u, v;
  uint32_t prim, dom;
};

// ONE routine for both query kinds over the warp-culled candidate domains (call warp_cull(..., true) first,
// all lanes).  A single instance keeps the persistent kernel's code small (instruction cache).
//   any == false: closest hit = lexicographic min of (t, domain, prim); returns out.t < inf.
//   any == true : returns true as soon as any resident entered domain occludes the ray.
template <bool COUNT, typename CandF>
__device__ __forceinline__ bool scene_trace(const DomainEntry* __restrict__ table, const DomSmem& D, const CandF& cand,
                                            vec3 o, vec3 d, const bool any, SceneHit& out,
                                            uint32_t* cnt /*[3]: domain tests, nodes, tris*/) {
  uint64_t list[MAXLIST];
  uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
  out.t = INFINITY; out.prim = 0; out.dom = 0xFFFFFFFFu; out.u = out.v = 0.f;
  const RayCtx r = make_ray_ctx(o, d, SPRAY_RAY_EPS);
  bool have_after = false, more = false, done = false;
  uint64_t after = 0;
  do {
    const int n = domain_list(D, cand, o, d, have_after, after, list, &more);
    for (int k = 0; k < n && !done; ++k) {
      const float tdom = sortable_float((uint32_t)(list[k] >> 32));
      // front-to-back early out (an optimisation only; slack keeps it conservative)
      if (out.t < INFINITY && tdom > out.t * 1.0001f + 1e-4f) { done = true; break; }
      const uint32_t id = (uint32_t)list[k];
      const DomainEntry* e = table + id;
      // Seed with the best t so far (prim = max): the traversal then admits hits with t < best, or
      // t == best (exact tie, resolved below to the lowest (domain, prim)).
      Closest cc; cc.t = out.t; cc.u = cc.v = 0.f; cc.prim = 0xFFFFFFFFu;
      if (COUNT) cnt[0]++;
      const bool occ = bvh8_trace(NodeLoaderG{e->nodes}, TriLoaderG{e->tris}, r, any, cc, sx, sy,
                                  COUNT ? cnt + 1 : nullptr, COUNT ? cnt + 2 : nullptr);
      if (occ) return true;
      if (cc.prim != 0xFFFFFFFFu && (cc.t < out.t || (cc.t == out.t && id < out.dom))) {
        out.t = cc.t; out.u = cc.u; out.v = cc.v; out.prim = cc.prim; out.dom = id;
      }
    }
    if (n > 0) { after = list[n - 1]; have_after = true; }
  } while (more && !done);
  return !any && out.t < INFINITY;
}

template <bool COUNT>
__device__ __forceinline__ void scene_closest(const DomainEntry* __restrict__ table, const DomSmem& D, vec3 o, vec3 d,
                                              SceneHit& out, uint32_t* cnt) {
  scene_trace<COUNT>(table, D, WarpCand{D.cand}, o, d, false, out, cnt);
}
template <bool COUNT>
__device__ __forceinline__ bool scene_occluded(const DomainEntry* __restrict__ table, const DomSmem& D, vec3 o, vec3 d,
                                               uint32_t* cnt) {
  SceneHit h;
  return scene_trace<COUNT>(table, D, WarpCand{D.cand}, o, d, true, h, cnt);
}

// TriMeshBuffer::updateIntersection (trimesh_buffer.cc:328-360) from the 48-byte attribute record
__device__ __forceinline__ void fetch_attributes(const DomainEntry* e, uint32_t prim, float u, float v,
                                                 uint32_t* color, vec3* Ns) {
  const float4* p = e->attr + 3 * (size_t)prim;
  const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  const float w = sub_rn(sub_rn(1.f, u), v);
  *color = interp_color(__float_as_uint(a.w), __float_as_uint(b.w), __float_as_uint(c.w), u, v, w);
  Ns->x = add_rn(add_rn(mul_rn(a.x, w), mul_rn(b.x, u)), mul_rn(c.x, v));
  Ns->y = add_rn(add_rn(mul_rn(a.y, w), mul_rn(b.y, u)), mul_rn(c.y, v));
  Ns->z = add_rn(add_rn(mul_rn(a.z, w), mul_rn(b.z, u)), mul_rn(c.z, v));
}

// ------------------------------------------------------------------------------------------------
// batched kernels.  No thread leaves before the warp-cooperative culling: out-of-range lanes stay as
// inactive participants.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
    domain_lists_kernel(const DomainEntry* __restrict__ table, int nd, const spraygpu_ray* __restrict__ rays, size_t n,
                        int cap, int32_t* __restrict__ counts, int32_t* __restrict__ ids, float* __restrict__ tmins) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(table, nd, smem);
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool valid = i < n;
  spraygpu_ray ry = {0, 0, 0, 0, 0, 0, 1, 0};
  if (valid) ry = rays[i];
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
  if (i >= n) return;
  const spraygpu_ray ry = rays[i];
  uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
  const RayCtx r = make_ray_ctx(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), ry.tmin);
  Closest c; c.t = INFINITY; c.u = c.v = 0.f; c.prim = 0xFFFFFFFFu;
  bvh8_closest(NodeLoaderG{nodes}, TriLoaderG{tris}, r, c, sx, sy, nullptr, nullptr);
  const bool hit = c.prim != 0xFFFFFFFFu;
  keys[i] = hit ? make_key(c.t, domain_id, c.prim) : SPRAY_KEY_MISS;
  if (uv) { uv[2 * i] = hit ? c.u : 0.f; uv[2 * i + 1] = hit ? c.v : 0.f; }
}

__global__ void __launch_bounds__(128)
    occluded_scene_kernel(const uint4* __restrict__ nodes, const float4* __restrict__ tris,
                          const spraygpu_ray* __restrict__ rays, size_t n, uint8_t* __restrict__ occ) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const spraygpu_ray ry = rays[i];
  uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
  const RayCtx r = make_ray_ctx(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), ry.tmin);
  occ[i] = bvh8_any(NodeLoaderG{nodes}, TriLoaderG{tris}, r, sx, sy, nullptr, nullptr) ? 1 : 0;
}

__global__ void __launch_bounds__(128)
    intersect_domains_kernel(const DomainEntry* __restrict__ table, int nd, const spraygpu_ray* __restrict__ rays,
                             size_t n, uint64_t* __restrict__ keys, float* __restrict__ uv, int merge) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(table, nd, smem);
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool valid = i < n;
  spraygpu_ray ry = {0, 0, 0, 0, 0, 0, 1, 0};
  if (valid) ry = rays[i];
  const vec3 o = mk3(ry.ox, ry.oy, ry.oz), d = mk3(ry.dx, ry.dy, ry.dz);
  warp_cull(D, valid, o, d, true);
  if (!valid) return;
  SceneHit h;
  scene_closest<false>(table, D, o, d, h, nullptr);
  const bool hit = h.t < INFINITY;
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
    occluded_domains_kernel(const DomainEntry* __restrict__ table, int nd, const spraygpu_ray* __restrict__ rays,
                            size_t n, uint8_t* __restrict__ occ, int merge) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(table, nd, smem);
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  // merge: rays already occluded elsewhere (or flagged 1 by the caller as "not a ray") are skipped
  const bool valid = i < n && !(merge && occ[i]);
  spraygpu_ray ry = {0, 0, 0, 0, 0, 0, 1, 0};
  if (valid) ry = rays[i];
  const vec3 o = mk3(ry.ox, ry.oy, ry.oz), d = mk3(ry.dx, ry.dy, ry.dz);
  warp_cull(D, valid, o, d, true);
  if (!valid) return;
  const bool oc = scene_occluded<false>(table, D, o, d, nullptr);
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
      const spraygpu_ray ry = rays[i];
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

__device__ __forceinline__ void eye_sample(const CameraDev& cam, spraygpu_tile tile, int spp, uint32_t x0,
                                           uint32_t y0, uint32_t sidx, uint32_t* samid, uint32_t* pixid, vec3* dir) {
  const uint32_t bufid = (uint32_t)spp * (y0 * (uint32_t)tile.w + x0) + sidx;
  const int x = tile.x + (int)x0, y = tile.y + (int)y0;
  float fx = (float)x, fy = (float)y;
  if (spp > 1) {  // genMultiEyes, ooc_tracer.inl:151-155
    uint32_t s = rng_init1(bufid);
    fx = add_rn((float)x, rng_float(s));
    fy = add_rn((float)y, rng_float(s));
  }
  *dir = camera_dir(cam, fx, fy);
  *samid = bufid;
  *pixid = (uint32_t)(y * cam.image_w + x);
}

__global__ void __launch_bounds__(256)
    eye_rays_kernel(CameraDev cam, spraygpu_tile tile, int spp, spraygpu_ray* __restrict__ rays) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t n = (size_t)tile.w * tile.h * spp;
  if (i >= n) return;
  const uint32_t sidx = (uint32_t)(i % spp);
  const uint32_t pix = (uint32_t)(i / spp);
  uint32_t samid, pixid; vec3 d;
  eye_sample(cam, tile, spp, pix % (uint32_t)tile.w, pix / (uint32_t)tile.w, sidx, &samid, &pixid, &d);
  spraygpu_ray r;
  r.ox = cam.pos[0]; r.oy = cam.pos[1]; r.oz = cam.pos[2]; r.tmin = SPRAY_RAY_EPS;
  r.dx = d.x; r.dy = d.y; r.dz = d.z; r.tmax = INFINITY;
  rays[samid] = r;
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
struct RenderParams {
  const DomainEntry* table;
  int nd;
  CameraDev cam;
  spraygpu_tile tile;
  spraygpu_shade shade;
  float* fb;                     // image_w * image_h * 4
  unsigned long long* counters;  // [0] work counter; [1..7] stats
  uint32_t total_slots;          // 32 * pixel blocks * spp
  uint32_t blocks_per_row;
};

struct Pending { vec3 o, d, w; int depth; };

#ifndef SPRAY_RENDER_MINBLOCKS
#define SPRAY_RENDER_MINBLOCKS 2
#endif

// ------------------------------------------------------------------------------------------------
// Per-warp ray queue (shared memory, SoA, LIFO).  Every ray that enters at least one resident domain
// — eye, bounce or shadow — is pushed with its exact domain mask; the warp pops 32 at a time, so all
// lanes traverse (ballot / prefix-sum compaction of the speculative ray stream).  Rays that enter no
// domain never occupy a lane.
// fields (QCAP entries each): 0-2 origin, 3-5 direction, 6-8 throughput (radiance) or radiance to add
// (shadow), 9 pixid, 10 samid, 11 depth << 1 | kind (0 radiance, 1 shadow), 12.. exact domain mask words
// ------------------------------------------------------------------------------------------------
#define QF_O 0
#define QF_D 3
#define QF_W 6
#define QF_PIX 9
#define QF_SAM 10
#define QF_DK 11
#define QF_MASK 12

struct RayQueue {
  uint32_t* q;        // this warp's fields, field f entry s at q[f * cap + s]
  uint32_t* scratch;  // this warp's provisional mask words, word w of lane l at scratch[w * 32 + l]
  int cap, nwords;
};

__host__ __device__ inline size_t queue_words_per_warp(int nd, int cap) {
  const int nw = (nd + 31) >> 5;
  return (size_t)(QF_MASK + nw) * cap + (size_t)nw * 32;
}

// Warp-cooperative push.  Returns true for lanes whose ray is live (emit) but enters NO resident domain:
// such a ray is resolved on the spot (radiance: miss, shadow: unoccluded).  `count` is warp-uniform.
__device__ __forceinline__ bool queue_push(const RayQueue& Q, const DomSmem& D, int& count, bool emit, vec3 o, vec3 d,
                                           vec3 w, uint32_t pixid, uint32_t samid, uint32_t depthkind) {
  const int lane = threadIdx.x & 31;
  warp_cull(D, emit, o, d, true);
  bool nonempty = false;
  if (emit) {
    for (int k = 0; k < Q.nwords; ++k) {
      const uint32_t m = exact_mask_word(D, k, D.cand[k], o, d);
      Q.scratch[k * 32 + lane] = m;
      nonempty = nonempty || (m != 0u);
    }
  }
  const bool go = emit && nonempty;
  const uint32_t b = __ballot_sync(0xffffffffu, go);
  if (go) {
    const int s = count + __popc(b & ((1u << lane) - 1u));
    uint32_t* q = Q.q + s;
    const int c = Q.cap;
    q[(QF_O + 0) * c] = __float_as_uint(o.x); q[(QF_O + 1) * c] = __float_as_uint(o.y); q[(QF_O + 2) * c] = __float_as_uint(o.z);
    q[(QF_D + 0) * c] = __float_as_uint(d.x); q[(QF_D + 1) * c] = __float_as_uint(d.y); q[(QF_D + 2) * c] = __float_as_uint(d.z);
    q[(QF_W + 0) * c] = __float_as_uint(w.x); q[(QF_W + 1) * c] = __float_as_uint(w.y); q[(QF_W + 2) * c] = __float_as_uint(w.z);
    q[QF_PIX * c] = pixid; q[QF_SAM * c] = samid; q[QF_DK * c] = depthkind;
    for (int k = 0; k < Q.nwords; ++k) q[(QF_MASK + k) * c] = Q.scratch[k * 32 + lane];
  }
  count += __popc(b);
  __syncwarp();
  return emit && !nonempty;
}

__device__ __forceinline__ void fb_add(float* fb, uint32_t pixid, float scale, vec3 Lr) {
  // TContext::retire -> HdrImage::add (ooc_tcontext.inl:123-135, image.h:99-109)
  float* px = fb + 4 * (size_t)pixid;
  atomicAdd(px + 0, mul_rn(scale, Lr.x));
  atomicAdd(px + 1, mul_rn(scale, Lr.y));
  atomicAdd(px + 2, mul_rn(scale, Lr.z));
}

// GENERAL == false is the lean variant for the common configuration (PT shader, ONE point light, every
// domain diffuse): area-light / AO sampling, delta materials and the inline shadow fallback are compiled
// out, which keeps the code the instruction cache has to hold small.  The launcher picks the variant.
template <bool GENERAL> struct QueueCap { static constexpr int value = GENERAL ? 128 : 96; };

template <bool COUNT, bool GENERAL>
__global__ void __launch_bounds__(256, SPRAY_RENDER_MINBLOCKS) render_tile_kernel(RenderParams P) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(P.table, P.nd, smem);
  constexpr int QCAP = QueueCap<GENERAL>::value;
  const int lane = threadIdx.x & 31;
  RayQueue Q;
  Q.cap = QCAP; Q.nwords = (P.nd + 31) >> 5;
  {
    uint32_t* qbase = (uint32_t*)((char*)smem + ((dom_smem_bytes(P.nd, 8) + 15) & ~(size_t)15));
    Q.q = qbase + (size_t)(threadIdx.x >> 5) * queue_words_per_warp(P.nd, QCAP);
    Q.scratch = Q.q + (size_t)(QF_MASK + Q.nwords) * QCAP;
  }
  const int spp = P.shade.pixel_samples;
  const float scale = (float)(1.0 / (double)spp);
  const vec3 ks = mk3(P.shade.ks[0], P.shade.ks[1], P.shade.ks[2]);
  const bool pt = GENERAL ? (P.shade.shader == SPRAYGPU_SHADER_PT) : true;
  const int nl = pt ? P.shade.nlights : 1;
  uint32_t cnt[3] = {0, 0, 0};
  uint32_t n_eye = 0, n_rad = 0, n_shadow = 0, n_hit = 0;
  int count = 0;      // rays in this warp's queue (warp-uniform)
  bool more = true;   // the tile still has unclaimed sample slots

  for (;;) {
    if (count < 32 && more) {
      // ---------------- refill: one 8x4 pixel block of eye rays (ooc_tracer.inl:84-172) ----------------
      unsigned long long base = 0;
      if (lane == 0) base = atomicAdd(P.counters, 32ull);
      base = __shfl_sync(0xffffffffu, base, 0);
      if (base >= P.total_slots) { more = false; continue; }
      const uint32_t slot = (uint32_t)base + lane;
      const uint32_t pix_slot = slot / (uint32_t)spp, sidx = slot % (uint32_t)spp;
      const uint32_t blk = pix_slot >> 5, within = pix_slot & 31u;
      const uint32_t x0 = (blk % P.blocks_per_row) * 8u + (within & 7u);
      const uint32_t y0 = (blk / P.blocks_per_row) * 4u + (within >> 3);
      const bool has_ray = slot < P.total_slots && x0 < (uint32_t)P.tile.w && y0 < (uint32_t)P.tile.h;
      uint32_t samid = 0, pixid = 0;
      vec3 dir = mk3(0.f, 0.f, 1.f);
      if (has_ray) {
        eye_sample(P.cam, P.tile, spp, x0, y0, sidx, &samid, &pixid, &dir);
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
    if (act) found = scene_trace<COUNT>(P.table, D, StridedCand{qe + QF_MASK * QCAP, QCAP}, ro, rd, shadow, h, cnt);
    // payload (read before any push overwrites the popped entries)
    const vec3 Lin = mk3(__uint_as_float(qe[(QF_W + 0) * QCAP]), __uint_as_float(qe[(QF_W + 1) * QCAP]),
                         __uint_as_float(qe[(QF_W + 2) * QCAP]));
    const uint32_t pixid = qe[QF_PIX * QCAP], samid = qe[QF_SAM * QCAP];
    __syncwarp();

    if (act && shadow && !found) fb_add(P.fb, pixid, scale, Lin);

    // ---------------- shade radiance hits (ooc_shader_pt.h:94-227, ooc_shader_ao.h:92-197) ----------------
    const bool hit = act && !shadow && found;
    const int next_depth = (int)(dk >> 1) + 1;
    vec3 pos = ro, kd = mk3(0, 0, 0), nff = mk3(0, 0, 1), wo = gneg(rd);
    float cos_theta_i = 0.f; bool entering = false; int mtl = SPRAYGPU_MTL_DIFFUSE; float eta0 = 1.f, eta1 = 1.f;
    if (hit) {
      ++n_hit;
      const DomainEntry* e = P.table + h.dom;
      uint32_t color; vec3 normal;
      fetch_attributes(e, h.prim, h.u, h.v, &color, &normal);
      pos = mk3(add_rn(mul_rn(rd.x, h.t), ro.x), add_rn(mul_rn(rd.y, h.t), ro.y), add_rn(mul_rn(rd.z, h.t), ro.z));
      kd = unpack_color(color);
      cos_theta_i = gdot(wo, normal);
      entering = cos_theta_i > 0.0f;
      nff = gnormalize(entering ? normal : gneg(normal));
      if (GENERAL) { mtl = e->mtl_type; eta0 = e->mtl[0]; eta1 = e->mtl[1]; }
    }
    const bool delta = mtl != SPRAYGPU_MTL_DIFFUSE;

    // ---------------- secondary rays: pushed first (the queue has room for them by construction) ----------------
    {
      bool em1 = false, em2 = false;
      vec3 d1 = mk3(0, 0, 1), w1 = mk3(0, 0, 0), d2 = mk3(0, 0, 1), w2 = mk3(0, 0, 0);
      if (hit && next_depth < P.shade.bounces) {
        const vec3 won = gnormalize(wo);
        if (GENERAL && delta) {
          if (cos_theta_i != 0.0f) {
            float ci = gclamp(cos_theta_i, -1.0f, 1.0f);
            const float abs_cos = fabsf(ci);
            if (!entering) ci = abs_cos;
            bool has_reflect = false, has_trans = false; float fr = 0.f; vec3 wt = mk3(0, 0, 0);
            if (mtl == SPRAYGPU_MTL_MIRROR) { fr = 1.0f; has_reflect = true; }
            else {
              const float etaI = entering ? eta0 : eta1, etaT = entering ? eta1 : eta0;
              if (mtl == SPRAYGPU_MTL_GLASS) {
                fr = fr_dielectric(ci, etaI, etaT, won, nff, &wt);
                if (fr == 1.0f) has_reflect = true;
                else if (fr == 0.0f) has_trans = true;
                else { has_reflect = true; has_trans = true; }
              } else {
                if (refract_dir(ci, etaI, etaT, won, nff, &wt)) { has_trans = true; fr = 0.0f; }
                else { has_reflect = true; fr = 1.0f; }
              }
            }
            // the reference CHECK-fails when glass yields both (ooc_shader_pt.h:199-201); we trace both
            if (has_reflect) {
              const vec3 wr = gadd(gneg(won), gscale(nff, mul_rn(2.0f, gdot(won, nff))));
              d1 = gnormalize(wr);
              w1 = gscale(Lin, div_rn(fr, abs_cos));
              em1 = has_positive(w1);
            }
            if (has_trans) {
              d2 = gnormalize(wt);
              w2 = gscale(Lin, div_rn(sub_rn(1.0f, fr), abs_cos));
              em2 = has_positive(w2);
            }
          }
        } else {
          uint32_t rs = rng_init1(samid * (uint32_t)next_depth);
          const float u1 = rng_float(rs), u2 = rng_float(rs);
          float pdf;
          cosine_hemisphere(u1, u2, nff, &d1, &pdf);
          const float costheta = gclamp(gdot(nff, d1), 0.0f, 1.0f);
          vec3 Lr = gscale(gscale(gmul(Lin, kd), SPRAY_ONE_OVER_PI), costheta);
          w1 = mk3(div_rn(Lr.x, pdf), div_rn(Lr.y, pdf), div_rn(Lr.z, pdf));
          em1 = has_positive(w1);
        }
      }
      if (em1) ++n_rad;
      if (__any_sync(0xffffffffu, em1))
        queue_push(Q, D, count, em1, pos, d1, w1, pixid, samid, (uint32_t)next_depth << 1);
      if (GENERAL) {
        if (em2) ++n_rad;
        if (__any_sync(0xffffffffu, em2))
          queue_push(Q, D, count, em2, pos, d2, w2, pixid, samid, (uint32_t)next_depth << 1);
      }
    }
    // ---------------- shadow rays: one light sample per lane per pass ----------------
    {
      uint32_t ls = rng_init1(samid * (uint32_t)next_depth);  // ShaderPt light_sampler (ooc_shader_pt.h:129-130)
      int l = 0, k = 0;
      bool more_lights = hit && !delta && nl > 0;
      while (__any_sync(0xffffffffu, more_lights)) {
        bool emit = false; vec3 wi = mk3(0, 0, 1), Lr = mk3(0, 0, 0);
        if (more_lights) {
          const spraygpu_light& L = P.shade.lights[pt ? l : 0];
          const bool area = GENERAL ? (pt ? (L.type == SPRAYGPU_LIGHT_DIFFUSE) : true) : false;
          const vec3 rad = mk3(L.radiance[0], L.radiance[1], L.radiance[2]);
          float costheta;
          if (GENERAL && area) {
            float pdf;
            if (pt) {
              const float u1 = rng_float(ls), u2 = rng_float(ls);
              cosine_hemisphere(u1, u2, nff, &wi, &pdf);
              if (pdf > 0.0f) {
                costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
                Lr = gscale(gmul(Lin, blinn_phong(costheta, kd, ks, P.shade.shininess, rad, wi, nff, wo)),
                            div_rn(1.0f, mul_rn(pdf, (float)P.shade.ao_samples)));
                emit = true;
              }
            } else {
              const float ao_weight = div_rn(1.0f, (float)P.shade.ao_samples);
              uint32_t as = rng_init1(pixid * (uint32_t)(k + 1));  // ooc_shader_ao.h:131
              const float u1 = rng_float(as), u2 = rng_float(as);
              cosine_hemisphere(u1, u2, nff, &wi, &pdf);
              costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
              Lr = gscale(gmul(Lin, kd), div_rn(mul_rn(mul_rn(SPRAY_ONE_OVER_PI, costheta), ao_weight), pdf));
              emit = true;
            }
          } else {
            wi = gnormalize(gsub(mk3(L.position[0], L.position[1], L.position[2]), pos));
            costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
            Lr = gscale(gmul(Lin, blinn_phong(costheta, kd, ks, P.shade.shininess, rad, wi, nff, wo)), 1.0f);
            emit = true;
          }
          emit = emit && has_positive(Lr);
          ++k;
          if (k >= (area ? P.shade.ao_samples : 1)) { k = 0; ++l; }
          more_lights = l < nl;
        }
        if (emit) ++n_shadow;
        if (!GENERAL || count + 32 <= QCAP) {
          // a shadow ray that enters no resident domain is unoccluded
          if (queue_push(Q, D, count, emit, pos, wi, Lr, pixid, samid, 1u)) fb_add(P.fb, pixid, scale, Lr);
        } else {
          // queue full (many light samples per hit): trace this pass in place
          warp_cull(D, emit, pos, wi, true);
          if (emit) {
            SceneHit hs;
            if (!scene_trace<COUNT>(P.table, D, WarpCand{D.cand}, pos, wi, true, hs, cnt)) fb_add(P.fb, pixid, scale, Lr);
          }
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
// host-side launchers (called from capi.cu)
// ------------------------------------------------------------------------------------------------
static inline int grid_for(size_t n, int t) { return (int)((n + t - 1) / t); }

static CameraDev to_camdev(const spraygpu_camera* c) {
  CameraDev d;
  for (int k = 0; k < 3; ++k) { d.pos[k] = c->pos[k]; d.lowerleft[k] = c->lowerleft[k]; d.wvec[k] = c->wvec[k]; d.hvec[k] = c->hvec[k]; }
  d.image_w = c->image_w; d.image_h = c->image_h;
  return d;
}

int launch_domain_lists(Device* dev, const spraygpu_ray* rays, size_t n, int cap, int32_t* counts, int32_t* ids,
                        float* tmins) {
  if (!n) return SPRAYGPU_OK;
  domain_lists_kernel<<<grid_for(n, 128), 128, dom_smem_bytes(dev->ndomains, 4), dev->stream>>>(
      dev->d_domains, dev->ndomains, rays, n, cap, counts, ids, tmins);
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
  intersect_domains_kernel<<<grid_for(n, 128), 128, dom_smem_bytes(dev->ndomains, 4), dev->stream>>>(
      dev->d_domains, dev->ndomains, rays, n, keys, uv, merge);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}
int launch_occluded_domains(Device* dev, const spraygpu_ray* rays, size_t n, uint8_t* occ, int merge) {
  if (!n) return SPRAYGPU_OK;
  occluded_domains_kernel<<<grid_for(n, 128), 128, dom_smem_bytes(dev->ndomains, 4), dev->stream>>>(
      dev->d_domains, dev->ndomains, rays, n, occ, merge);
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
int launch_eye_rays(Device* dev, const spraygpu_camera* cam, spraygpu_tile tile, int spp, spraygpu_ray* rays) {
  const size_t n = (size_t)tile.w * tile.h * spp;
  if (!n) return SPRAYGPU_OK;
  eye_rays_kernel<<<grid_for(n, 256), 256, 0, dev->stream>>>(to_camdev(cam), tile, spp, rays);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

int launch_render_tile(Device* dev, const spraygpu_camera* cam, const spraygpu_shade* shade, spraygpu_tile tile,
                       float* fb, bool counted) {
  RenderParams P;
  P.table = dev->d_domains; P.nd = dev->ndomains;
  P.cam = to_camdev(cam); P.tile = tile; P.shade = *shade; P.fb = fb;
  P.counters = dev->d_counters;
  P.blocks_per_row = (uint32_t)((tile.w + 7) / 8);
  const uint64_t slots = (uint64_t)P.blocks_per_row * (uint64_t)((tile.h + 3) / 4) * 32ull * (uint64_t)shade->pixel_samples;
  SG_CHECK(slots < 0xFFFFFFFFull, SPRAYGPU_ERR_LIMIT, "tile too large: %llu sample slots", (unsigned long long)slots);
  P.total_slots = (uint32_t)slots;
  if (!slots) return SPRAYGPU_OK;
  SG_CUDA(cudaMemsetAsync(dev->d_counters, 0, sizeof(unsigned long long) * 8, dev->stream));
  // lean variant: PT shader, ONE point light, all domains diffuse
  bool general = shade->shader != SPRAYGPU_SHADER_PT || shade->nlights != 1;
  for (int i = 0; i < shade->nlights; ++i) general = general || shade->lights[i].type != SPRAYGPU_LIGHT_POINT;
  for (int i = 0; i < dev->ndomains; ++i) general = general || dev->h_domains[i].mtl_type != SPRAYGPU_MTL_DIFFUSE;
  void (*kern)(RenderParams) =
      counted ? (general ? render_tile_kernel<true, true> : render_tile_kernel<true, false>)
              : (general ? render_tile_kernel<false, true> : render_tile_kernel<false, false>);
  const int qcap = general ? QueueCap<true>::value : QueueCap<false>::value;
  const size_t smem = ((dom_smem_bytes(dev->ndomains, 8) + 15) & ~(size_t)15) +
                      8 * queue_words_per_warp(dev->ndomains, qcap) * sizeof(uint32_t);
  SG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, smem));
  if (per_sm < 1) per_sm = 1;
  // persistent grid: SM count x resident CTAs per SM
  int grid = dev->sm_count * per_sm;
  const uint64_t warps_needed = (slots + 31) / 32;
  if ((uint64_t)grid * 8 > warps_needed) grid = (int)((warps_needed + 7) / 8);
  kern<<<grid, 256, smem, dev->stream>>>(P);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

}  // namespace spraygpu
