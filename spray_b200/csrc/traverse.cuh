// traverse.cuh — device-side traversal building blocks shared by trace.cu and insitu.cu: global-memory
// node / triangle loaders, the warp-cooperative domain culling (beam test + ballot), the per-lane exact
// domain list, and scene_trace (closest / any hit over the entered resident domains).
#pragma once
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

// ------------------------------------------------------------------------------------------------
// Warp-synchronous traversal.
//
// ncu on the per-lane while-while walk above (profiles/r01c): 4.7 active threads per warp instruction on
// incoherent bounce rays — once lanes split inside the nested loops (domain loop x node loop x triangle
// loop, with early returns) they never re-merge.  Here ALL 32 lanes of the warp run ONE flat loop and meet
// at a __syncwarp at its head; a lane's own progress is a small state machine:
//     idle  -> pop the traversal stack, or pull its next entered domain (front to back)
//     node  -> fetch + test one 80-byte BVH8 node        (only when it has no pending triangles)
//     tri   -> fetch + test ONE pending triangle
// so every pass of the loop executes the node test once and the triangle test once for the whole warp.
// The arithmetic per (ray, node) and (ray, triangle) is unchanged, and the closest hit is the
// lexicographic min of (t, domain, prim) — independent of visiting order — so results are bit-identical
// to the per-lane walk.
//
// DomainSource: next(best_t, nodes, tris, dom) -> false when the lane has no further domain to walk.
// ------------------------------------------------------------------------------------------------
struct OneDomainSource {   // seam 2: a single committed scene
  const uint4* nodes; const float4* tris; uint32_t id; bool used;
  __device__ __forceinline__ bool next(float, const uint4*& n, const float4*& t, uint32_t& dom) {
    if (used) return false;
    used = true; n = nodes; t = tris; dom = id;
    return true;
  }
};

template <typename CandF>
struct ListDomainSource {  // the ray's entered resident domains, front to back by (tmin, id)
  const DomainEntry* table; const DomSmem& D; const CandF& cand; vec3 o, d;
  uint64_t list[MAXLIST];
  int n, k; bool more;
  __device__ __forceinline__ ListDomainSource(const DomainEntry* table_, const DomSmem& D_, const CandF& cand_, bool act,
                                              vec3 o_, vec3 d_)
      : table(table_), D(D_), cand(cand_), o(o_), d(d_), n(0), k(0), more(false) {
    if (act) n = domain_list(D, cand, o, d, false, 0ull, list, &more);
  }
  __device__ __forceinline__ bool next(float best_t, const uint4*& nn, const float4*& tt, uint32_t& dom) {
    for (;;) {
      if (k < n) {
        const uint64_t key = list[k];
        const float tdom = sortable_float((uint32_t)(key >> 32));
        // front-to-back early out (an optimisation only; the slack keeps it conservative)
        if (best_t < INFINITY && tdom > best_t * 1.0001f + 1e-4f) return false;
        const DomainEntry* e = table + (uint32_t)key;
        nn = e->nodes; tt = e->tris; dom = (uint32_t)key; ++k;
        return true;
      }
      if (!more || n == 0) return false;
      const uint64_t after = list[n - 1];   // more than MAXLIST entered domains: next chunk
      n = domain_list(D, cand, o, d, true, after, list, &more);
      k = 0;
    }
  }
};

// Call with ALL 32 lanes of the warp (act = this lane carries a ray).  Returns: any == false -> out.t < inf;
// any == true -> occluded.
template <bool COUNT, typename Source>
__device__ __forceinline__ bool trace_warp(Source& src, const bool act, const RayCtx& r, const bool any, SceneHit& out,
                                           uint32_t* cnt /*[3]: domain tests, nodes, tris*/) {
  uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
  out.t = INFINITY; out.prim = 0xFFFFFFFFu; out.dom = 0xFFFFFFFFu; out.u = out.v = 0.f;
  bool alive = act, found = false;
  uint32_t ngx = 0, ngy = 0, tgx = 0, tgy = 0, dom = 0;
  const uint4* nodes = nullptr;
  const float4* tris = nullptr;
  int sp = 0;
  for (;;) {
    __syncwarp();
    if (alive && ngy <= 0x00FFFFFFu && tgy == 0u) {   // idle: pop, or pull the next domain
      if (sp > 0) { --sp; ngx = sx[sp]; ngy = sy[sp]; }
      else if (src.next(out.t, nodes, tris, dom)) { ngx = 0u; ngy = 0x80000000u; if (COUNT) cnt[0]++; }
      else alive = false;
    }
    if (!__any_sync(0xffffffffu, alive)) break;
    if (alive && tgy == 0u) {                          // node step (ngy has an inner child here)
      const uint32_t imask = ngy & 0xffu;
      const int bit = 31 - __clz((int)ngy);
      ngy &= ~(1u << bit);
      if (ngy > 0x00FFFFFFu && sp < SPRAY_STACK_SIZE) { sx[sp] = ngx; sy[sp] = ngy; ++sp; }
      const uint32_t slot = (uint32_t)(bit - 24) ^ r.octinv;
      const uint32_t rel = (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot));
      u4 q0, q1, q2, q3, q4;
      NodeLoaderG{nodes}(ngx + rel, q0, q1, q2, q3, q4);
      if (COUNT) cnt[1]++;
      const uint32_t hm = node8_hits(q0, q1, q2, q3, q4, r, out.t);
      ngx = q1.x; tgx = q1.y;
      ngy = (hm & 0xFF000000u) | (q0.w >> 24);
      tgy = hm & 0x00FFFFFFu;
    }
    __syncwarp();
    if (alive && tgy) {                                // triangle step: ONE pending triangle
      const int ti = __ffs((int)tgy) - 1;
      tgy &= tgy - 1;
      vec3 v0, e1, e2; uint32_t prim;
      TriLoaderG{tris}(tgx + (uint32_t)ti, v0, e1, e2, prim);
      if (COUNT) cnt[2]++;
      float T, U, V, ad;
      if (tri_test_raw(r.o, r.d, v0, e1, e2, r.tnear, &T, &U, &V, &ad)) {
        if (any) { found = true; alive = false; }
        else {
          const float t = div_rn(T, ad);
          if (t < out.t || (t == out.t && (dom < out.dom || (dom == out.dom && prim < out.prim)))) {
            out.t = t; out.u = div_rn(U, ad); out.v = div_rn(V, ad); out.prim = prim; out.dom = dom;
          }
        }
      }
    }
  }
  return any ? found : (out.t < INFINITY);
}

// closest / any hit over the ray's entered resident domains; all 32 lanes call.
template <bool COUNT, typename CandF>
__device__ __forceinline__ bool scene_trace_warp(const DomainEntry* __restrict__ table, const DomSmem& D,
                                                 const CandF& cand, const bool act, vec3 o, vec3 d, const bool any,
                                                 SceneHit& out, uint32_t* cnt) {
  const RayCtx r = make_ray_ctx(o, d, SPRAY_RAY_EPS);
  ListDomainSource<CandF> src(table, D, cand, act, o, d);
  return trace_warp<COUNT>(src, act, r, any, out, cnt);
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

}  // namespace spraygpu
