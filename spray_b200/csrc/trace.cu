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

// per-lane exact test on the candidates: keeps the MAXLIST nearest entries with key > after.
__device__ __forceinline__ int domain_list(const DomSmem& D, vec3 o, vec3 d, bool have_after, uint64_t after,
                                           uint64_t* list, bool* more) {
  int cnt = 0;
  bool overflow = false;
  const int nwords = (D.nd + 31) >> 5;
  for (int w = 0; w < nwords; ++w) {
    uint32_t m = D.cand[w];
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

struct SceneHit {
  float t, u, v;
  uint32_t prim, dom;
};

// closest hit over the warp-culled candidate domains; call warp_cull(..., true) first (all lanes).
template <bool COUNT>
__device__ __forceinline__ void scene_closest(const DomainEntry* __restrict__ table, const DomSmem& D, vec3 o, vec3 d,
                                              SceneHit& out, uint32_t* cnt /*[3]: domain tests, nodes, tris*/) {
  uint64_t list[MAXLIST];
  uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
  out.t = INFINITY; out.prim = 0; out.dom = 0xFFFFFFFFu; out.u = out.v = 0.f;
  const RayCtx r = make_ray_ctx(o, d, SPRAY_RAY_EPS);
  bool have_after = false, more = false, done = false;
  uint64_t after = 0;
  do {
    const int n = domain_list(D, o, d, have_after, after, list, &more);
    for (int k = 0; k < n && !done; ++k) {
      const float tdom = sortable_float((uint32_t)(list[k] >> 32));
      // front-to-back early out (an optimisation only; slack keeps it conservative)
      if (out.t < INFINITY && tdom > out.t * 1.0001f + 1e-4f) { done = true; break; }
      const uint32_t id = (uint32_t)list[k];
      const DomainEntry* e = table + id;
      // Seed with the best t so far (prim = max): the traversal then admits hits with t < best, or
      // t == best (exact tie, resolved below to the lowest (domain, prim)).
      Closest cc; cc.t = out.t; cc.u = cc.v = 0.f; cc.prim = 0xFFFFFFFFu;
      bvh8_closest(NodeLoaderG{e->nodes}, TriLoaderG{e->tris}, r, cc, sx, sy, COUNT ? cnt + 1 : nullptr,
                   COUNT ? cnt + 2 : nullptr);
      if (COUNT) cnt[0]++;
      if (cc.prim != 0xFFFFFFFFu && (cc.t < out.t || (cc.t == out.t && id < out.dom))) {
        out.t = cc.t; out.u = cc.u; out.v = cc.v; out.prim = cc.prim; out.dom = id;
      }
    }
    if (n > 0) { after = list[n - 1]; have_after = true; }
  } while (more && !done);
}

template <bool COUNT>
__device__ __forceinline__ bool scene_occluded(const DomainEntry* __restrict__ table, const DomSmem& D, vec3 o, vec3 d,
                                               uint32_t* cnt) {
  uint64_t list[MAXLIST];
  uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
  const RayCtx r = make_ray_ctx(o, d, SPRAY_RAY_EPS);
  bool have_after = false, more = false;
  uint64_t after = 0;
  do {
    const int n = domain_list(D, o, d, have_after, after, list, &more);
    for (int k = 0; k < n; ++k) {
      const DomainEntry* e = table + (uint32_t)list[k];
      if (COUNT) cnt[0]++;
      if (bvh8_any(NodeLoaderG{e->nodes}, TriLoaderG{e->tris}, r, sx, sy, COUNT ? cnt + 1 : nullptr,
                   COUNT ? cnt + 2 : nullptr))
        return true;
    }
    if (n > 0) { after = list[n - 1]; have_after = true; }
  } while (more);
  return false;
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
    const int c = domain_list(D, o, d, have_after, after, list, &more);
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
#define SPRAY_RENDER_MINBLOCKS 3
#endif
template <bool COUNT>
__global__ void __launch_bounds__(256, SPRAY_RENDER_MINBLOCKS) render_tile_kernel(RenderParams P) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(P.table, P.nd, smem);

  const int lane = threadIdx.x & 31;
  const int spp = P.shade.pixel_samples;
  const float scale = (float)(1.0 / (double)spp);
  const vec3 ks = mk3(P.shade.ks[0], P.shade.ks[1], P.shade.ks[2]);
  const bool pt = P.shade.shader == SPRAYGPU_SHADER_PT;
  const int nl = pt ? P.shade.nlights : 1;
  uint32_t cnt[3] = {0, 0, 0};
  uint32_t n_eye = 0, n_rad = 0, n_shadow = 0, n_hit = 0;

  for (;;) {
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(P.counters, 32ull);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= P.total_slots) break;
    const uint32_t slot = (uint32_t)base + lane;
    // slot -> (pixel block, lane in block, sample)
    const uint32_t pix_slot = slot / (uint32_t)spp, sidx = slot % (uint32_t)spp;
    const uint32_t blk = pix_slot >> 5, within = pix_slot & 31u;
    const uint32_t x0 = (blk % P.blocks_per_row) * 8u + (within & 7u);
    const uint32_t y0 = (blk / P.blocks_per_row) * 4u + (within >> 3);
    bool has_ray = slot < P.total_slots && x0 < (uint32_t)P.tile.w && y0 < (uint32_t)P.tile.h;

    uint32_t samid = 0, pixid = 0;
    Pending cur;
    cur.o = mk3(P.cam.pos[0], P.cam.pos[1], P.cam.pos[2]);
    cur.d = mk3(0.f, 0.f, 1.f); cur.w = mk3(1.f, 1.f, 1.f); cur.depth = 0;
    if (has_ray) {
      eye_sample(P.cam, P.tile, spp, x0, y0, sidx, &samid, &pixid, &cur.d);
      ++n_eye;
    }
    Pending stack[MAX_PENDING];  // second rays of glass hits (reflection + transmission); rarely used
    int nstack = 0;

    while (__any_sync(0xffffffffu, has_ray)) {
      // ---------------- closest hit ----------------
      warp_cull(D, has_ray, cur.o, cur.d, true);
      SceneHit h; h.t = INFINITY; h.dom = 0; h.prim = 0; h.u = h.v = 0.f;
      if (has_ray) { scene_closest<COUNT>(P.table, D, cur.o, cur.d, h, cnt); ++n_rad; }
      const bool hit = has_ray && h.t < INFINITY;
      // ---------------- shading context ----------------
      vec3 pos = cur.o, kd = mk3(0, 0, 0), nff = mk3(0, 0, 1), wo = gneg(cur.d);
      float cos_theta_i = 0.f; bool entering = false; int mtl = SPRAYGPU_MTL_DIFFUSE; float eta0 = 1.f, eta1 = 1.f;
      if (hit) {
        ++n_hit;
        const DomainEntry* e = P.table + h.dom;
        uint32_t color; vec3 normal;
        fetch_attributes(e, h.prim, h.u, h.v, &color, &normal);
        pos = mk3(add_rn(mul_rn(cur.d.x, h.t), cur.o.x), add_rn(mul_rn(cur.d.y, h.t), cur.o.y),
                  add_rn(mul_rn(cur.d.z, h.t), cur.o.z));
        kd = unpack_color(color);
        cos_theta_i = gdot(wo, normal);
        entering = cos_theta_i > 0.0f;
        nff = gnormalize(entering ? normal : gneg(normal));
        mtl = e->mtl_type; eta0 = e->mtl[0]; eta1 = e->mtl[1];
      }
      const bool delta = mtl != SPRAYGPU_MTL_DIFFUSE;
      const int next_depth = cur.depth + 1;
      const vec3 Lin = cur.w;
      // ---------------- shadow rays: one light sample per warp pass ----------------
      {
        uint32_t ls = rng_init1(samid * (uint32_t)next_depth);  // ShaderPt light_sampler (ooc_shader_pt.h:129-130)
        const float ao_weight = div_rn(1.0f, (float)P.shade.ao_samples);
        int l = 0, k = 0;
        bool more_lights = hit && !delta && nl > 0;
        while (__any_sync(0xffffffffu, more_lights)) {
          bool emit = false; vec3 wi = mk3(0, 0, 1), Lr = mk3(0, 0, 0);
          if (more_lights) {
            const spraygpu_light& L = P.shade.lights[pt ? l : 0];
            const bool area = pt ? (L.type == SPRAYGPU_LIGHT_DIFFUSE) : true;
            const vec3 rad = mk3(L.radiance[0], L.radiance[1], L.radiance[2]);
            float pdf, costheta;
            if (pt) {
              if (area) {
                const float u1 = rng_float(ls), u2 = rng_float(ls);
                cosine_hemisphere(u1, u2, nff, &wi, &pdf);
                if (pdf > 0.0f) {
                  costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
                  Lr = gscale(gmul(Lin, blinn_phong(costheta, kd, ks, P.shade.shininess, rad, wi, nff, wo)),
                              div_rn(1.0f, mul_rn(pdf, (float)P.shade.ao_samples)));
                  emit = true;
                }
              } else {
                wi = gnormalize(gsub(mk3(L.position[0], L.position[1], L.position[2]), pos));
                costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
                Lr = gscale(gmul(Lin, blinn_phong(costheta, kd, ks, P.shade.shininess, rad, wi, nff, wo)), 1.0f);
                emit = true;
              }
            } else {
              uint32_t as = rng_init1(pixid * (uint32_t)(k + 1));  // ooc_shader_ao.h:131
              const float u1 = rng_float(as), u2 = rng_float(as);
              cosine_hemisphere(u1, u2, nff, &wi, &pdf);
              costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
              Lr = gscale(gmul(Lin, kd), div_rn(mul_rn(mul_rn(SPRAY_ONE_OVER_PI, costheta), ao_weight), pdf));
              emit = true;
            }
            emit = emit && has_positive(Lr);
            // advance (l, k)
            ++k;
            if (k >= (area ? P.shade.ao_samples : 1)) { k = 0; ++l; }
            more_lights = l < nl;
          }
          warp_cull(D, emit, pos, wi, true);
          if (emit) {
            ++n_shadow;
            if (!scene_occluded<COUNT>(P.table, D, pos, wi, cnt)) {
              // TContext::retire -> HdrImage::add (ooc_tcontext.inl:123-135, image.h:99-109)
              float* px = P.fb + 4 * (size_t)pixid;
              atomicAdd(px + 0, mul_rn(scale, Lr.x));
              atomicAdd(px + 1, mul_rn(scale, Lr.y));
              atomicAdd(px + 2, mul_rn(scale, Lr.z));
            }
          }
        }
      }
      // ---------------- next ray of this lane: bounce, or a pending second ray ----------------
      bool next = false;
      if (hit && next_depth < P.shade.bounces) {
        wo = gnormalize(wo);
        vec3 wi, Lr; float pdf, costheta;
        if (delta) {
          if (cos_theta_i != 0.0f) {
            cos_theta_i = gclamp(cos_theta_i, -1.0f, 1.0f);
            const float abs_cos = fabsf(cos_theta_i);
            if (!entering) cos_theta_i = abs_cos;
            bool has_reflect = false, has_trans = false; float fr = 0.f; vec3 wt = mk3(0, 0, 0);
            if (mtl == SPRAYGPU_MTL_MIRROR) { fr = 1.0f; has_reflect = true; }
            else {
              const float etaI = entering ? eta0 : eta1, etaT = entering ? eta1 : eta0;
              if (mtl == SPRAYGPU_MTL_GLASS) {
                fr = fr_dielectric(cos_theta_i, etaI, etaT, wo, nff, &wt);
                if (fr == 1.0f) has_reflect = true;
                else if (fr == 0.0f) has_trans = true;
                else { has_reflect = true; has_trans = true; }
              } else {
                if (refract_dir(cos_theta_i, etaI, etaT, wo, nff, &wt)) { has_trans = true; fr = 0.0f; }
                else { has_reflect = true; fr = 1.0f; }
              }
            }
            // the reference CHECK-fails when glass yields both (ooc_shader_pt.h:199-201); we trace both:
            // reflection first, transmission parked on the lane's pending stack
            Pending refl, trans; bool ok_r = false, ok_t = false;
            if (has_trans) {
              wi = gnormalize(wt);
              Lr = gscale(Lin, div_rn(sub_rn(1.0f, fr), abs_cos));
              if (has_positive(Lr)) { trans.o = pos; trans.d = wi; trans.w = Lr; trans.depth = next_depth; ok_t = true; }
            }
            if (has_reflect) {
              const vec3 wr = gadd(gneg(wo), gscale(nff, mul_rn(2.0f, gdot(wo, nff))));
              wi = gnormalize(wr);
              Lr = gscale(Lin, div_rn(fr, abs_cos));
              if (has_positive(Lr)) { refl.o = pos; refl.d = wi; refl.w = Lr; refl.depth = next_depth; ok_r = true; }
            }
            if (ok_r) { cur = refl; next = true; if (ok_t && nstack < MAX_PENDING) stack[nstack++] = trans; }
            else if (ok_t) { cur = trans; next = true; }
          }
        } else {
          uint32_t rs = rng_init1(samid * (uint32_t)next_depth);
          const float u1 = rng_float(rs), u2 = rng_float(rs);
          cosine_hemisphere(u1, u2, nff, &wi, &pdf);
          costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
          Lr = gscale(gscale(gmul(Lin, kd), SPRAY_ONE_OVER_PI), costheta);
          Lr = mk3(div_rn(Lr.x, pdf), div_rn(Lr.y, pdf), div_rn(Lr.z, pdf));
          if (has_positive(Lr)) { cur.o = pos; cur.d = wi; cur.w = Lr; cur.depth = next_depth; next = true; }
        }
      }
      if (!next && has_ray && nstack > 0) { cur = stack[--nstack]; next = true; }
      has_ray = next;
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
  const size_t smem = dom_smem_bytes(dev->ndomains, 8);
  int per_sm = 0;
  if (counted) {
    SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, render_tile_kernel<true>, 256, smem));
  } else {
    SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, render_tile_kernel<false>, 256, smem));
  }
  if (per_sm < 1) per_sm = 1;
  // persistent grid: SM count x resident CTAs per SM
  int grid = dev->sm_count * per_sm;
  const uint64_t warps_needed = (slots + 31) / 32;
  if ((uint64_t)grid * 8 > warps_needed) grid = (int)((warps_needed + 7) / 8);
  if (counted) render_tile_kernel<true><<<grid, 256, smem, dev->stream>>>(P);
  else render_tile_kernel<false><<<grid, 256, smem, dev->stream>>>(P);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

}  // namespace spraygpu
