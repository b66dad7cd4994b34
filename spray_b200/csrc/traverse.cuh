// traverse.cuh — device-side traversal building blocks shared by trace.cu and insitu.cu: global-memory
// node / triangle loaders, the warp-cooperative domain culling (beam test + ballot), the per-lane exact
// domain list, and scene_trace (closest / any hit over the entered resident domains).
#pragma once
#include "internal.cuh"

namespace spraygpu {

#define MAXLIST 16
#define MAX_PENDING 8

// ray records move as two 16-byte transactions (spraygpu_ray arrays are 16-byte aligned: include/spray_gpu.h)
__device__ __forceinline__ spraygpu_ray load_ray(const spraygpu_ray* p) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  spraygpu_ray r;
  r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
  return r;
}
__device__ __forceinline__ void store_ray(spraygpu_ray* p, const spraygpu_ray& r) {
  reinterpret_cast<float4*>(p)[0] = make_float4(r.ox, r.oy, r.oz, r.tmin);
  reinterpret_cast<float4*>(p)[1] = make_float4(r.dx, r.dy, r.dz, r.tmax);
}

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
  uint32_t* cand;           // this warp's CAND_WORDS candidate words (domain_lists parity kernel only)
  const TlasNode* tlas;     // top-level BVH4 over the resident domains (16-byte aligned)
  int nd, tlas_n;
};

// layout: tlas nodes | boxes | resident | [candidate words]
__host__ __device__ inline size_t dom_smem_bytes(int nd, int tlas_n, int cand_warps) {
  return (size_t)tlas_n * sizeof(TlasNode) + (size_t)nd * 6 * sizeof(float) + (((size_t)nd + 15) & ~(size_t)15) +
         (size_t)cand_warps * CAND_WORDS * 4;
}

__device__ __forceinline__ DomSmem load_domain_smem(const DomainEntry* __restrict__ table, int nd,
                                                    const TlasNode* __restrict__ tlas, int tlas_n, float* smem) {
  uint4* tl = (uint4*)smem;
  float* boxes = (float*)(tl + 7 * tlas_n);
  uint8_t* resident = (uint8_t*)(boxes + 6 * nd);
  uint32_t* cand = (uint32_t*)(resident + ((nd + 15) & ~15));
  for (int i = threadIdx.x; i < 7 * tlas_n; i += blockDim.x) tl[i] = ((const uint4*)tlas)[i];
  for (int i = threadIdx.x; i < nd; i += blockDim.x) {
#pragma unroll
    for (int k = 0; k < 6; ++k) boxes[6 * i + k] = table[i].box[k];
    resident[i] = (uint8_t)table[i].resident;
  }
  __syncthreads();
  DomSmem d;
  d.boxes = boxes; d.resident = resident; d.cand = cand + (threadIdx.x >> 5) * CAND_WORDS; d.nd = nd;
  d.tlas = (const TlasNode*)tl; d.tlas_n = tlas_n;
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
  // one step of "find my next domain": 0 = no further domain, 1 = domain found, 2 = progressed, call again
  __device__ __forceinline__ int step(const RayCtx&, float, const uint4*& n, const float4*& t, uint32_t& dom) {
    if (used) return 0;
    used = true; n = nodes; t = tris; dom = id;
    return 1;
  }
};

// ------------------------------------------------------------------------------------------------
// TlasSource — the ray's entered resident domains found LAZILY by walking the top-level BVH4 (shared
// memory) front to back.  ncu/A-B on the flat scheme (exact slab test of every candidate box at push time):
// incoherent bounce rays tested all 64 boxes (~1.4 k instructions per ray) although 3 of 4 rays end in the
// first domain they enter.  Here a ray only opens the top-level nodes in front of its current best hit.
//   inner boxes : conservative float slab test (same (plane - o) * idir form as the exact test, which is
//                 monotone in the box, plus a relative tolerance)
//   leaf (domain): the reference's exact slab test (aabb.h:139-169) decides entry, as before
// Order only affects speed: the closest hit is the lexicographic min of (t, domain, prim).
// ------------------------------------------------------------------------------------------------
#define SPRAY_TLAS_STACK 24
#ifndef SPRAY_SLAB_HOIST
#define SPRAY_SLAB_HOIST 0
#endif

__device__ __forceinline__ bool tlas_box_hit(float lox, float loy, float loz, float hix, float hiy, float hiz, const RayCtx& r,
                                             float* tmin_o) {
  const bool nx = r.idir.x < 0.0f, ny = r.idir.y < 0.0f, nz = r.idir.z < 0.0f;
  const float t0x = ((nx ? hix : lox) - r.o.x) * r.idir.x, t1x = ((nx ? lox : hix) - r.o.x) * r.idir.x + r.far_bias.x;
  const float t0y = ((ny ? hiy : loy) - r.o.y) * r.idir.y, t1y = ((ny ? loy : hiy) - r.o.y) * r.idir.y + r.far_bias.y;
  const float t0z = ((nz ? hiz : loz) - r.o.z) * r.idir.z, t1z = ((nz ? loz : hiz) - r.o.z) * r.idir.z + r.far_bias.z;
  const float cmin = fmaxf(fmaxf(t0x, t0y), t0z), cmax = fminf(fminf(t1x, t1y), t1z);
  const float tol = 1e-5f * (fabsf(cmin) + fabsf(cmax)) + 1e-30f;
  *tmin_o = cmin - tol;
  return (cmin <= cmax + tol) && (cmax >= -tol);
}

// The stacks are objects of their own, NOT members of the state structs: a store to st[sp] inside the struct may
// alias every scalar member as far as ptxas can tell, and the first build reloaded sp / ngx / ngy / the node
// pointers from local memory after each push (LDL/STL pairs all over the hot loop in the SASS).
struct TlasStack { float t[SPRAY_TLAS_STACK]; int r[SPRAY_TLAS_STACK]; };
struct TlasSource {
  const DomainEntry* table; const DomSmem& D;
  float* const st;
  int* const sr;
  int sp;
  __device__ __forceinline__ TlasSource(const DomainEntry* table_, const DomSmem& D_, TlasStack& S)
      : table(table_), D(D_), st(S.t), sr(S.r), sp(0) {}
  __device__ __forceinline__ void reset(bool act) {
    sp = 0;
    if (act && D.tlas_n > 0) { st[0] = -INFINITY; sr[0] = 0; sp = 1; }
  }
  // ONE top-level step: pops one entry; 0 = stack empty, 1 = entered a domain, 2 = progressed
  __device__ __forceinline__ int step(const RayCtx& r, float best_t, const uint4*& nn, const float4*& tt, uint32_t& dom) {
    {
      // entries behind the current best hit are dropped without a step of their own (cheap)
      while (sp > 0 && best_t < INFINITY && st[sp - 1] > best_t * 1.0001f + 1e-4f) --sp;
      if (sp == 0) return 0;
      --sp;
      const int ref = sr[sp];
      if (ref < 0) {
        const uint32_t id = (uint32_t)~ref;
        float tmin;
#if SPRAY_SLAB_HOIST
        // A/B knob (r02c, not yet measured on its own): the reference divides 1/d in every box test; idir already
        // holds those quotients except on clamped axes (tests/test_hostemu.py checks the equivalence bit for bit)
        float ix = r.idir.x, iy = r.idir.y, iz = r.idir.z;
        if (!(fminf(fminf(fabsf(r.d.x), fabsf(r.d.y)), fabsf(r.d.z)) > SPRAY_IDIR_CLAMP)) {
          ix = exact_inv(r.d.x, ix); iy = exact_inv(r.d.y, iy); iz = exact_inv(r.d.z, iz);
        }
        if (!domain_slab_inv(D.boxes + 6 * id, r.o, ix, iy, iz, SPRAY_RAY_EPS, INFINITY, &tmin)) return 2;
#else
        if (!domain_slab(D.boxes + 6 * id, r.o, r.d, SPRAY_RAY_EPS, INFINITY, &tmin)) return 2;
#endif
        const DomainEntry* e = table + id;
        nn = e->nodes; tt = e->tris; dom = id;
        return 1;
      }
      const float4* n4 = (const float4*)(D.tlas + ref);
      const float4 lx = n4[0], ly = n4[1], lz = n4[2], hx = n4[3], hy = n4[4], hz = n4[5];
      const int4 ch = *(const int4*)(n4 + 6);
      float t[4]; int c[4];
      bool h0 = tlas_box_hit(lx.x, ly.x, lz.x, hx.x, hy.x, hz.x, r, &t[0]);
      bool h1 = tlas_box_hit(lx.y, ly.y, lz.y, hx.y, hy.y, hz.y, r, &t[1]);
      bool h2 = tlas_box_hit(lx.z, ly.z, lz.z, hx.z, hy.z, hz.z, r, &t[2]);
      bool h3 = tlas_box_hit(lx.w, ly.w, lz.w, hx.w, hy.w, hz.w, r, &t[3]);
      c[0] = ch.x; c[1] = ch.y; c[2] = ch.z; c[3] = ch.w;
      if (!h0) t[0] = INFINITY;
      if (!h1) t[1] = INFINITY;
      if (!h2) t[2] = INFINITY;
      if (!h3) t[3] = INFINITY;
      // sort the four children far -> near (5 compare-exchanges), push the hits: nearest ends on top
#define SPRAY_CSWAP(a, b) { if (t[a] < t[b]) { const float tf = t[a]; t[a] = t[b]; t[b] = tf; const int ti = c[a]; c[a] = c[b]; c[b] = ti; } }
      SPRAY_CSWAP(0, 1) SPRAY_CSWAP(2, 3) SPRAY_CSWAP(0, 2) SPRAY_CSWAP(1, 3) SPRAY_CSWAP(1, 2)
#undef SPRAY_CSWAP
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (t[k] < INFINITY) {
          // upload_domain_table refuses top-level trees whose worst-case stack (3 per level + 1) exceeds SPRAY_TLAS_STACK
          if (sp >= SPRAY_TLAS_STACK) SPRAY_STACK_OVERFLOW();
          st[sp] = t[k]; sr[sp] = c[k]; ++sp;
        }
    }
    return 2;
  }
};

// conservative "does this ray enter any resident domain at all" (the top-level root's children): rays that
// miss the whole scene never take a queue slot
__device__ __forceinline__ bool tlas_any(const DomSmem& D, const RayCtx& r) {
  if (D.tlas_n == 0) return false;
  const float4* n4 = (const float4*)D.tlas;
  const float4 lx = n4[0], ly = n4[1], lz = n4[2], hx = n4[3], hy = n4[4], hz = n4[5];
  float t;
  return tlas_box_hit(lx.x, ly.x, lz.x, hx.x, hy.x, hz.x, r, &t) || tlas_box_hit(lx.y, ly.y, lz.y, hx.y, hy.y, hz.y, r, &t) ||
         tlas_box_hit(lx.z, ly.z, lz.z, hx.z, hy.z, hz.z, r, &t) || tlas_box_hit(lx.w, ly.w, lz.w, hx.w, hy.w, hz.w, r, &t);
}

#ifndef SPRAY_TRI_PER_PASS
#define SPRAY_TRI_PER_PASS 8
#endif
// ------------------------------------------------------------------------------------------------
// Warp-cooperative triangle stage (SPRAY_COOP_TRI, north_star's "triangle blocks staged in shared memory").
// ncu on the per-lane form (profiles/r01y_c3_render_tile_lines.txt): every triangle-test line runs at ~4 of 32 lanes —
// few lanes have pending triangles in a given pass, each works its own list off serially (up to 8 dependent 48-byte
// fetches), and the other lanes wait at the next __syncwarp.  Here the warp pools its pending (owner lane, triangle)
// pairs in a shared-memory list (ballot-free prefix sum), then ALL lanes take one pair each: fetch the 48-byte
// record, read the owner's ray by shuffle, run the same S1 test, and merge hits into the owner's slot with a 64-bit
// shared-memory atomicMin on (t bits | primID) — the packed-key update of north_star.  The winner writes (u, v).
// Same arithmetic per (ray, triangle) and the same (t, domain, prim) tie rule, so hit records stay bit-identical.
//   1: node step, then the pooled triangle stage (triangles of this pass's node are tested in the same pass)
//   2: software-pipelined — the triangles found by the PREVIOUS node step are fetched and tested while this pass's
//      node fetch is in flight (one dependent memory round trip per pass instead of two)
// ------------------------------------------------------------------------------------------------
#ifndef SPRAY_COOP_TRI
#define SPRAY_COOP_TRI 0
#endif
#ifndef SPRAY_COOP_K
#define SPRAY_COOP_K 4          // triangles one lane may contribute per pass
#endif
#define SPRAY_MAX_CTA_WARPS 4          // every kernel that traverses runs 128-thread CTAs (static_assert in trace.cu)
struct CoopScratch {
  uint32_t pair[32 * SPRAY_COOP_K];     // owner lane << 27 | triangle index (leaf order, < 2^24)
  unsigned long long best[32];          // per owner: min over this pass of (t bits << 32 | primID)
  float2 uv[32];                        // (u, v) of that minimum
  uint32_t anyflag;                     // any-hit mode: bit per owner
};
// One lane's traversal state and its three step kinds.
#if SPRAY_POP_CULL
struct LaneStack { uint4 e[SPRAY_STACK_SIZE]; };   // (child base, pending mask, distance bytes of slots 0..3, 4..7)
#else
struct LaneStack { uint32_t x[SPRAY_STACK_SIZE], y[SPRAY_STACK_SIZE]; };   // separate object: see TlasStack
#endif
struct LaneTrav {
#if SPRAY_POP_CULL
  uint4* const stk;
  uint32_t gd0, gd1;   // distance bytes of the current group (core.cuh, SPRAY_POP_CULL)
  uint32_t oct;        // the ray's octinv, for the pruning done inside want()
#else
  uint32_t* const sx;
  uint32_t* const sy;
#endif
  uint32_t ngx, ngy, tgx, tgy, dom;
  const uint4* nodes;
  const float4* tris;
  int sp;
  bool alive, found;
  SceneHit out;

#if SPRAY_POP_CULL
  __device__ __forceinline__ explicit LaneTrav(LaneStack& S) : stk(S.e), gd0(0u), gd1(0u), oct(0u) {}
#else
  __device__ __forceinline__ explicit LaneTrav(LaneStack& S) : sx(S.x), sy(S.y) {}
#endif
  __device__ __forceinline__ void begin(bool act) {
    alive = act; found = false;
    ngx = ngy = tgx = tgy = 0u; dom = 0u; sp = 0;
    nodes = nullptr; tris = nullptr;
    out.t = INFINITY; out.prim = 0xFFFFFFFFu; out.dom = 0xFFFFFFFFu; out.u = out.v = 0.f;
  }
  // What this lane needs next.
  enum { WANT_NONE = 0, WANT_SOURCE = 1, WANT_NODE = 2, WANT_TRI = 3 };
  __device__ __forceinline__ int want() {
    if (!alive) return WANT_NONE;
    if (tgy) return WANT_TRI;
#if SPRAY_POP_CULL
    while (ngy <= 0x00FFFFFFu && sp > 0) {   // pop; children that the best hit has overtaken since the push are dropped here
      --sp;
      const uint4 e = stk[sp];
      ngx = e.x; gd0 = e.z; gd1 = e.w;
      ngy = prune_group(e.y, gd0, gd1, out.t, oct);
    }
#else
    if (ngy <= 0x00FFFFFFu && sp > 0) { --sp; ngx = sx[sp]; ngy = sy[sp]; }   // popping costs nothing: do it now
#endif
    return ngy > 0x00FFFFFFu ? WANT_NODE : WANT_SOURCE;
  }
  // one step of the domain source (top-level BVH walk); a lane whose source is exhausted is finished
  template <bool COUNT, typename Source>
  __device__ __forceinline__ void source_step(Source& src, const RayCtx& r, uint32_t* cnt) {
    int rc;
    do { rc = src.step(r, out.t, nodes, tris, dom); } while (rc == 2);   // A/B: one top-level step per pass was 3-4 % slower
    if (rc == 1) {
      ngx = 0u; ngy = 0x80000000u; if (COUNT) cnt[0]++;
#if SPRAY_POP_CULL
      gd0 = gd1 = 0u;
#endif
    }
    else if (rc == 0) alive = false;
  }
  // One pass of the warp-synchronous loop: every lane takes at most one step of each kind, in the order
  // source -> node -> triangle, so a lane that just entered a domain tests its root in the same pass.
  // (A/B on B200: letting the warp vote for ONE kind of step per pass was 8-12 % slower — the minority lanes
  // stall instead of riding along.)
  template <bool COUNT, typename Source>
  __device__ __forceinline__ void pass(Source& src, const RayCtx& r, const bool any, uint32_t* cnt) {
#if SPRAY_POP_CULL
    oct = r.octinv;
#endif
    int w = want();
    if (w == WANT_SOURCE) { source_step<COUNT>(src, r, cnt); w = want(); }
    __syncwarp();
    if (w == WANT_NODE) node_step<COUNT>(r, cnt);
    __syncwarp();
#if SPRAY_TRI_PER_PASS <= 1
    if (alive && tgy) tri_step<COUNT>(r, any, cnt);
#else
    // a triangle test costs ~35 thread instructions, a node step ~450 (ncu r01j): let a lane work off several
    // pending triangles per pass so that it is back in the node test sooner.  A/B on B200 (profiles/r01l_ab_tri_per_pass.txt):
    // 1 -> 8 triangles per pass = C2 5862 -> 6776 Mrays/s, C3 928 -> 1041
#pragma unroll 1
    for (int k = 0; k < SPRAY_TRI_PER_PASS && alive && tgy; ++k) tri_step<COUNT>(r, any, cnt);
#endif
  }
#if SPRAY_COOP_TRI
  // what a lane WITHOUT pending triangles needs next (pops the stack on the way)
  __device__ __forceinline__ int want_no_tri() {
    if (!alive) return WANT_NONE;
#if SPRAY_POP_CULL
    while (ngy <= 0x00FFFFFFu && sp > 0) {
      --sp;
      const uint4 e = stk[sp];
      ngx = e.x; gd0 = e.z; gd1 = e.w;
      ngy = prune_group(e.y, gd0, gd1, out.t, oct);
    }
#else
    if (ngy <= 0x00FFFFFFu && sp > 0) { --sp; ngx = sx[sp]; ngy = sy[sp]; }
#endif
    return ngy > 0x00FFFFFFu ? WANT_NODE : WANT_SOURCE;
  }
  // pool this lane's pending triangles (at most SPRAY_COOP_K) into the warp's pair list; returns the warp total
  template <bool COUNT>
  __device__ __forceinline__ int coop_contribute(CoopScratch& S, uint32_t* cnt) {
    const int lane = threadIdx.x & 31;
    int nt = alive ? min(__popc(tgy), SPRAY_COOP_K) : 0;
    int incl = nt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += o;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return 0;
    int at = incl - nt;
    for (int k = 0; k < nt; ++k) {
      const int ti = __ffs((int)tgy) - 1;
      tgy &= tgy - 1;
      S.pair[at + k] = ((uint32_t)lane << 27) | (tgx + (uint32_t)ti);
    }
    if (COUNT) cnt[2] += (uint32_t)nt;
    S.best[lane] = ~0ull;
    if (lane == 0) S.anyflag = 0u;
    __syncwarp();
    return total;
  }
  // all lanes: one pair each per round.  ptris / pdom: the contributing lane's triangle array and domain AT CONTRIBUTION
  // time (a source step later in the pass may already have moved the lane to its next domain).
  __device__ __forceinline__ void coop_test(CoopScratch& S, int total, const RayCtx& r, const bool any, const float4* ptris) {
    const int lane = threadIdx.x & 31;
    for (int base = 0; base < total; base += 32) {
      const int idx = base + lane;
      const bool valid = idx < total;
      const uint32_t pr = valid ? S.pair[idx] : ((uint32_t)lane << 27);
      const int owner = (int)(pr >> 27);
      const vec3 O = mk3(__shfl_sync(0xffffffffu, r.o.x, owner), __shfl_sync(0xffffffffu, r.o.y, owner), __shfl_sync(0xffffffffu, r.o.z, owner));
      const vec3 D = mk3(__shfl_sync(0xffffffffu, r.d.x, owner), __shfl_sync(0xffffffffu, r.d.y, owner), __shfl_sync(0xffffffffu, r.d.z, owner));
      const float tn = __shfl_sync(0xffffffffu, r.tnear, owner);
      const unsigned long long tp = __shfl_sync(0xffffffffu, (unsigned long long)ptris, owner);
      const bool oany = __shfl_sync(0xffffffffu, (int)any, owner) != 0;
      bool hit = false;
      float T = 0.f, U = 0.f, V = 0.f, ad = 1.f;
      unsigned long long key = ~0ull;
      if (valid) {
        vec3 v0, e1, e2; uint32_t prim;
        TriLoaderG{(const float4*)tp}(pr & 0x07FFFFFFu, v0, e1, e2, prim);
        hit = tri_test_raw(O, D, v0, e1, e2, tn, &T, &U, &V, &ad);
        if (hit) {
          if (oany) atomicOr(&S.anyflag, 1u << owner);
          else {
            key = ((unsigned long long)__float_as_uint(div_rn(T, ad)) << 32) | (unsigned long long)prim;
            atomicMin(&S.best[owner], key);
          }
        }
      }
      __syncwarp();
      if (hit && !oany && S.best[owner] == key) S.uv[owner] = make_float2(div_rn(U, ad), div_rn(V, ad));
      __syncwarp();
    }
  }
  // owner side: fold the pooled result into this lane's closest hit / any-hit flag
  __device__ __forceinline__ void coop_merge(CoopScratch& S, const RayCtx& r, const bool any, uint32_t pdom) {
    const int lane = threadIdx.x & 31;
    if (any) {
      if ((S.anyflag >> lane) & 1u) { found = true; alive = false; }
      return;
    }
    const unsigned long long key = S.best[lane];
    if (key == ~0ull) return;
    const float t = __uint_as_float((uint32_t)(key >> 32));
    const uint32_t prim = (uint32_t)key;
    if (t < out.t || (t == out.t && (pdom < out.dom || (pdom == out.dom && prim < out.prim)))) {
      const float2 uv = S.uv[lane];
      out.t = t; out.u = uv.x; out.v = uv.y; out.prim = prim; out.dom = pdom;
#if SPRAY_POP_CULL
      ngy = prune_group(ngy, gd0, gd1, out.t, r.octinv);
#endif
    }
  }
  template <bool COUNT, typename Source>
  __device__ __forceinline__ void pass_coop(Source& src, const RayCtx& r, const bool any, uint32_t* cnt, CoopScratch& S) {
#if SPRAY_POP_CULL
    oct = r.octinv;
#endif
#if SPRAY_COOP_TRI == 1
    // node step first; its triangles are pooled and tested in this same pass
    int w = (alive && !tgy) ? want_no_tri() : WANT_NONE;
    if (w == WANT_SOURCE) { source_step<COUNT>(src, r, cnt); w = (alive && !tgy) ? want_no_tri() : WANT_NONE; }
    __syncwarp();
    if (w == WANT_NODE) node_step<COUNT>(r, cnt);
    __syncwarp();
    const float4* ptris = tris; const uint32_t pdom = dom;
    const int total = coop_contribute<COUNT>(S, cnt);
    if (total) { coop_test(S, total, r, any, ptris); coop_merge(S, r, any, pdom); }
#else
    // pipelined: pool the triangles of the PREVIOUS node step, issue this pass's node fetch, test the triangles while
    // the node is in flight, then do the node arithmetic
    const float4* ptris = tris; const uint32_t pdom = dom;
    const int total = coop_contribute<COUNT>(S, cnt);
    int w = (alive && !tgy) ? want_no_tri() : WANT_NONE;
    if (w == WANT_SOURCE) { source_step<COUNT>(src, r, cnt); w = (alive && !tgy) ? want_no_tri() : WANT_NONE; }
    __syncwarp();
    u4 q0, q1, q2, q3, q4;
    const bool donode = w == WANT_NODE;
    if (donode) node_fetch<COUNT>(r, cnt, q0, q1, q2, q3, q4);
    if (total) coop_test(S, total, r, any, ptris);
    if (donode) node_math(r, q0, q1, q2, q3, q4);
    if (total) coop_merge(S, r, any, pdom);
#endif
  }
  // node_step split in two so that other loads can be issued between the fetch and the arithmetic
  template <bool COUNT>
  __device__ __forceinline__ void node_fetch(const RayCtx& r, uint32_t* cnt, u4& q0, u4& q1, u4& q2, u4& q3, u4& q4) {
    const uint32_t imask = ngy & 0xffu;
    const int bit = 31 - __clz((int)ngy);
    ngy &= ~(1u << bit);
    if (ngy > 0x00FFFFFFu && sp >= SPRAY_STACK_SIZE) SPRAY_STACK_OVERFLOW();
#if SPRAY_POP_CULL
    if (ngy > 0x00FFFFFFu) { stk[sp] = make_uint4(ngx, ngy, gd0, gd1); ++sp; }
#else
    if (ngy > 0x00FFFFFFu) { sx[sp] = ngx; sy[sp] = ngy; ++sp; }
#endif
    const uint32_t slot = (uint32_t)(bit - 24) ^ r.octinv;
    const uint32_t rel = (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot));
    NodeLoaderG{nodes}(ngx + rel, q0, q1, q2, q3, q4);
    if (COUNT) cnt[1]++;
  }
  __device__ __forceinline__ void node_math(const RayCtx& r, const u4& q0, const u4& q1, const u4& q2, const u4& q3, const u4& q4) {
#if SPRAY_POP_CULL
    const uint32_t hm = node8_hits(q0, q1, q2, q3, q4, r, out.t, &gd0, &gd1);
#else
    const uint32_t hm = node8_hits(q0, q1, q2, q3, q4, r, out.t);
#endif
    ngx = q1.x; tgx = q1.y;
    ngy = (hm & 0xFF000000u) | (q0.w >> 24);
    tgy = hm & 0x00FFFFFFu;
  }
#endif  // SPRAY_COOP_TRI
  // fetch + test one 80-byte node (only when no triangle is pending; ngy has an inner child here)
  template <bool COUNT>
  __device__ __forceinline__ void node_step(const RayCtx& r, uint32_t* cnt) {
    {
      const uint32_t imask = ngy & 0xffu;
      const int bit = 31 - __clz((int)ngy);
      ngy &= ~(1u << bit);
      // commit refuses BVH8 trees deeper than the stack (core.cuh, SPRAY_STACK_SIZE): the trap is unreachable
      if (ngy > 0x00FFFFFFu && sp >= SPRAY_STACK_SIZE) SPRAY_STACK_OVERFLOW();
#if SPRAY_POP_CULL
      if (ngy > 0x00FFFFFFu) { stk[sp] = make_uint4(ngx, ngy, gd0, gd1); ++sp; }
#else
      if (ngy > 0x00FFFFFFu) { sx[sp] = ngx; sy[sp] = ngy; ++sp; }
#endif
      const uint32_t slot = (uint32_t)(bit - 24) ^ r.octinv;
      const uint32_t rel = (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot));
      u4 q0, q1, q2, q3, q4;
      NodeLoaderG{nodes}(ngx + rel, q0, q1, q2, q3, q4);
      if (COUNT) cnt[1]++;
#if SPRAY_POP_CULL
      const uint32_t hm = node8_hits(q0, q1, q2, q3, q4, r, out.t, &gd0, &gd1);
#else
      const uint32_t hm = node8_hits(q0, q1, q2, q3, q4, r, out.t);
#endif
      ngx = q1.x; tgx = q1.y;
      ngy = (hm & 0xFF000000u) | (q0.w >> 24);
      tgy = hm & 0x00FFFFFFu;
    }
  }
  // fetch + test ONE pending triangle
  template <bool COUNT>
  __device__ __forceinline__ void tri_step(const RayCtx& r, const bool any, uint32_t* cnt) {
    {
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
#if SPRAY_POP_CULL
            ngy = prune_group(ngy, gd0, gd1, out.t, r.octinv);
#endif
          }
        }
      }
    }
  }
};

// Call with ALL 32 lanes of the warp (act = this lane carries a ray).  Returns: any == false -> out.t < inf;
// any == true -> occluded.
// tbound: a hit is only wanted if (t, domain, prim) is lexicographically below (tbound, ~0, ~0) — the t of a hit already
// known elsewhere (another GPU's domains, ray forwarding); everything behind it is culled like a local best hit.
template <bool COUNT, typename Source>
__device__ __forceinline__ bool trace_warp(Source& src, const bool act, const RayCtx& r, const bool any, SceneHit& out,
                                           uint32_t* cnt /*[3]: domain tests, nodes, tris*/, const float tbound = INFINITY) {
  LaneStack stack;
  LaneTrav T(stack);
  T.begin(act);
  T.out.t = tbound;
#if SPRAY_COOP_TRI
  __shared__ CoopScratch coop_scratch[SPRAY_MAX_CTA_WARPS];
  CoopScratch& S = coop_scratch[threadIdx.x >> 5];
#endif
  for (;;) {
    __syncwarp();
    if (!__any_sync(0xffffffffu, T.alive)) break;
#if SPRAY_COOP_TRI
    T.pass_coop<COUNT>(src, r, any, cnt, S);
#else
    T.pass<COUNT>(src, r, any, cnt);
#endif
  }
  out = T.out;
  return any ? T.found : (out.prim != 0xFFFFFFFFu);
}

// closest / any hit over the ray's entered resident domains; all 32 lanes call.
template <bool COUNT>
__device__ __forceinline__ bool scene_trace_warp(const DomainEntry* __restrict__ table, const DomSmem& D, const bool act,
                                                 vec3 o, vec3 d, const bool any, SceneHit& out, uint32_t* cnt,
                                                 const float tbound = INFINITY) {
  const RayCtx r = make_ray_ctx(o, d, SPRAY_RAY_EPS);
  TlasStack tstack;
  TlasSource src(table, D, tstack);
  src.reset(act);
  return trace_warp<COUNT>(src, act, r, any, out, cnt, tbound);
}

}  // namespace spraygpu
