// core.cuh — building blocks of the B200 ray-tracing path, written as __host__ __device__ functions so
// that the exact same logic can be exercised by the host-emulation harness (tests/hostemu) in the
// GPU-less build container.  The product only ever runs them inside CUDA kernels (trace.cu, build.cu).
//
// FP32 contract (DESIGN.md "FP32 contract"):
//   * R-ops (mul_rn/add_rn/sub_rn/div_rn/sqrt_rn) round once and are never contracted: they restate the
//     reference's glm/C++ arithmetic (glm 0.9.9.3 dot = (x*x'+y*y')+z*z', normalize = v*(1/sqrt(dot))).
//   * The ray/triangle test (S1) uses explicit fused multiply-adds in a fixed order.
//   * Box tests (BVH8 nodes) are free-form FP32: only conservativeness matters there.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define SPRAY_HD __host__ __device__ __forceinline__
#else
#define SPRAY_HD inline
#endif

namespace spraygpu {

// ---------------------------------------------------------------------------------------------
// exactly-rounded scalar ops
// ---------------------------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
SPRAY_HD float mul_rn(float a, float b) { return __fmul_rn(a, b); }
SPRAY_HD float add_rn(float a, float b) { return __fadd_rn(a, b); }
SPRAY_HD float sub_rn(float a, float b) { return __fsub_rn(a, b); }
SPRAY_HD float div_rn(float a, float b) { return __fdiv_rn(a, b); }
SPRAY_HD float sqrt_rn(float a) { return __fsqrt_rn(a); }
SPRAY_HD float fma_rn(float a, float b, float c) { return __fmaf_rn(a, b, c); }
SPRAY_HD uint32_t f2u(float f) { return __float_as_uint(f); }
SPRAY_HD float u2f(uint32_t u) { return __uint_as_float(u); }
SPRAY_HD int clz32(uint32_t x) { return __clz((int)x); }
SPRAY_HD int popc32(uint32_t x) { return __popc(x); }
SPRAY_HD int ffs32(uint32_t x) { return __ffs((int)x); }
#else
// host emulation: compiled with -ffp-contract=off, so plain operators round once
SPRAY_HD float mul_rn(float a, float b) { return a * b; }
SPRAY_HD float add_rn(float a, float b) { return a + b; }
SPRAY_HD float sub_rn(float a, float b) { return a - b; }
SPRAY_HD float div_rn(float a, float b) { return a / b; }
SPRAY_HD float sqrt_rn(float a) { return sqrtf(a); }
SPRAY_HD float fma_rn(float a, float b, float c) { return fmaf(a, b, c); }
SPRAY_HD uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
SPRAY_HD float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
SPRAY_HD int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
SPRAY_HD int popc32(uint32_t x) { return __builtin_popcount(x); }
SPRAY_HD int ffs32(uint32_t x) { return __builtin_ffs((int)x); }
#endif

struct vec3 { float x, y, z; };
SPRAY_HD vec3 mk3(float x, float y, float z) { vec3 r; r.x = x; r.y = y; r.z = z; return r; }
// reference-order (glm) helpers
SPRAY_HD vec3 gadd(vec3 a, vec3 b) { return mk3(add_rn(a.x, b.x), add_rn(a.y, b.y), add_rn(a.z, b.z)); }
SPRAY_HD vec3 gsub(vec3 a, vec3 b) { return mk3(sub_rn(a.x, b.x), sub_rn(a.y, b.y), sub_rn(a.z, b.z)); }
SPRAY_HD vec3 gmul(vec3 a, vec3 b) { return mk3(mul_rn(a.x, b.x), mul_rn(a.y, b.y), mul_rn(a.z, b.z)); }
SPRAY_HD vec3 gscale(vec3 a, float s) { return mk3(mul_rn(a.x, s), mul_rn(a.y, s), mul_rn(a.z, s)); }
SPRAY_HD vec3 gneg(vec3 a) { return mk3(-a.x, -a.y, -a.z); }
SPRAY_HD float gdot(vec3 a, vec3 b) { return add_rn(add_rn(mul_rn(a.x, b.x), mul_rn(a.y, b.y)), mul_rn(a.z, b.z)); }
SPRAY_HD vec3 gcross(vec3 x, vec3 y) {
  return mk3(sub_rn(mul_rn(x.y, y.z), mul_rn(y.y, x.z)), sub_rn(mul_rn(x.z, y.x), mul_rn(y.z, x.x)),
             sub_rn(mul_rn(x.x, y.y), mul_rn(y.x, x.y)));
}
SPRAY_HD vec3 gnormalize(vec3 v) { return gscale(v, div_rn(1.0f, sqrt_rn(gdot(v, v)))); }
SPRAY_HD float gclamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// fused helpers of the triangle test (contract S1)
SPRAY_HD float edot(vec3 a, vec3 b) { return fma_rn(a.x, b.x, fma_rn(a.y, b.y, mul_rn(a.z, b.z))); }
SPRAY_HD vec3 ecross(vec3 a, vec3 b) {
  return mk3(fma_rn(a.y, b.z, -mul_rn(a.z, b.y)), fma_rn(a.z, b.x, -mul_rn(a.x, b.z)),
             fma_rn(a.x, b.y, -mul_rn(a.y, b.x)));
}

#define SPRAY_RAY_EPS 0.001f              /* SPRAY_RAY_EPSILON, reference src/render/spray.h:46 */
#define SPRAY_ONE_OVER_PI 0.3183098861837907f /* src/render/spray.h:50 */
#define SPRAY_KEY_MISS 0xFFFFFFFFFFFFFFFFull
#define SPRAY_DOMAIN_BITS 9
#define SPRAY_PRIM_BITS 23

SPRAY_HD uint64_t make_key(float t, uint32_t dom, uint32_t prim) {
  return ((uint64_t)f2u(t) << 32) | ((uint64_t)dom << SPRAY_PRIM_BITS) | (uint64_t)prim;
}
SPRAY_HD float key_t(uint64_t k) { return u2f((uint32_t)(k >> 32)); }
SPRAY_HD uint32_t key_domain(uint64_t k) { return (uint32_t)((k >> SPRAY_PRIM_BITS) & 0x1ff); }
SPRAY_HD uint32_t key_prim(uint64_t k) { return (uint32_t)(k & 0x7fffff); }

// ---------------------------------------------------------------------------------------------
// RNG — reference deps/embree/random_sampler.h:34-108 (MurmurHash3 seed + LCG; integer exact)
// ---------------------------------------------------------------------------------------------
SPRAY_HD uint32_t murmur_mix(uint32_t hash, uint32_t k) {
  k *= 0xcc9e2d51u;
  k = (k << 15) | (k >> 17);
  k *= 0x1b873593u;
  hash ^= k;
  return ((hash << 13) | (hash >> 19)) * 5u + 0xe6546b64u;
}
SPRAY_HD uint32_t murmur_fin(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
SPRAY_HD uint32_t rng_init1(uint32_t id) { return murmur_fin(murmur_mix(0u, id)); }
SPRAY_HD uint32_t rng_init2(uint32_t a, uint32_t b) { return murmur_fin(murmur_mix(murmur_mix(0u, a), b)); }
SPRAY_HD float rng_float(uint32_t& s) {
  s = s * 1664525u + 1013904223u;
  return mul_rn((float)(int)(s >> 1), 4.656612873077392578125e-10f);
}

// Morton — reference src/render/morton.h:33-52 (10 bits / axis)
SPRAY_HD uint32_t expand_bits10(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
SPRAY_HD uint32_t morton30(float x, float y, float z) {
  x = fminf(fmaxf(mul_rn(x, 1024.0f), 0.0f), 1023.0f);
  y = fminf(fmaxf(mul_rn(y, 1024.0f), 0.0f), 1023.0f);
  z = fminf(fmaxf(mul_rn(z, 1024.0f), 0.0f), 1023.0f);
  return expand_bits10((uint32_t)x) * 4 + expand_bits10((uint32_t)y) * 2 + expand_bits10((uint32_t)z);
}

// ---------------------------------------------------------------------------------------------
// Domain slab test — reference src/render/aabb.h:139-169, behaviour-identical (early-out order, NaN
// handling through the comparison directions, tmin not clamped).  b = {lo.xyz, hi.xyz}.
// ---------------------------------------------------------------------------------------------
// domain_slab_inv takes the three reciprocals 1/d (IEEE, as the reference computes them per test) so that a caller
// testing many boxes with one ray divides once.
SPRAY_HD bool domain_slab_inv(const float* b, vec3 o, float ix, float iy, float iz, float t0, float t1, float* tmin_o) {
  int sx = ix < 0.0f, sy = iy < 0.0f, sz = iz < 0.0f;
  float tmin = mul_rn(sub_rn(b[sx * 3 + 0], o.x), ix);
  float tmax = mul_rn(sub_rn(b[(1 - sx) * 3 + 0], o.x), ix);
  float tymin = mul_rn(sub_rn(b[sy * 3 + 1], o.y), iy);
  float tymax = mul_rn(sub_rn(b[(1 - sy) * 3 + 1], o.y), iy);
  if ((tmin > tymax) || (tymin > tmax)) return false;
  if (tymin > tmin) tmin = tymin;
  if (tymax < tmax) tmax = tymax;
  float tzmin = mul_rn(sub_rn(b[sz * 3 + 2], o.z), iz);
  float tzmax = mul_rn(sub_rn(b[(1 - sz) * 3 + 2], o.z), iz);
  if ((tmin > tzmax) || (tzmin > tmax)) return false;
  if (tzmin > tmin) tmin = tzmin;
  if (tzmax < tmax) tmax = tzmax;
  *tmin_o = tmin;
  return (tmin < t1) && (tmax > t0);
}
SPRAY_HD bool domain_slab(const float* b, vec3 o, vec3 d, float t0, float t1, float* tmin_o) {
  return domain_slab_inv(b, o, div_rn(1.0f, d.x), div_rn(1.0f, d.y), div_rn(1.0f, d.z), t0, t1, tmin_o);
}

// order-preserving float -> uint (handles negative entry distances: origin inside the box)
SPRAY_HD uint32_t float_sortable(float f) {
  uint32_t u = f2u(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
SPRAY_HD float sortable_float(uint32_t s) {
  uint32_t u = (s & 0x80000000u) ? (s & 0x7fffffffu) : ~s;
  return u2f(u);
}

// ---------------------------------------------------------------------------------------------
// Triangle record: 3 x float4 (48 B): (v0.xyz, primID bits) (e1.xyz, -) (e2.xyz, -),
// e1 = v0 - v1, e2 = v2 - v0 computed once with sub_rn (same values the per-ray subtraction gives).
// Contract S1: Moeller-Trumbore in the form used by Embree's triangle-v intersector [unverified, the
// dependency is not in the container]: no culling, inclusive edges, u,v relative to v1,v2, t > tnear.
// ---------------------------------------------------------------------------------------------
struct TriHit { float t, u, v; };
// pass/fail part (no divisions): scaled T, U, V and |den|
SPRAY_HD bool tri_test_raw(vec3 O, vec3 D, vec3 v0, vec3 e1, vec3 e2, float tnear, float* T, float* U, float* V,
                           float* absDen) {
  vec3 Ng = ecross(e2, e1);
  vec3 C = gsub(v0, O);
  vec3 R = ecross(C, D);
  float den = edot(Ng, D);
  float ad = fabsf(den);
  uint32_t sgn = f2u(den) & 0x80000000u;
  float u = u2f(f2u(edot(R, e2)) ^ sgn);
  float v = u2f(f2u(edot(R, e1)) ^ sgn);
  if (!(den != 0.0f) || !(u >= 0.0f) || !(v >= 0.0f) || !(add_rn(u, v) <= ad)) return false;
  float t = u2f(f2u(edot(Ng, C)) ^ sgn);
  if (!(t > mul_rn(ad, tnear))) return false;
  *T = t; *U = u; *V = v; *absDen = ad;
  return true;
}
SPRAY_HD bool tri_test(vec3 O, vec3 D, vec3 v0, vec3 e1, vec3 e2, float tnear, TriHit* h) {
  float T, U, V, ad;
  if (!tri_test_raw(O, D, v0, e1, e2, tnear, &T, &U, &V, &ad)) return false;
  h->t = div_rn(T, ad);
  h->u = div_rn(U, ad);
  h->v = div_rn(V, ad);
  return true;
}

// ---------------------------------------------------------------------------------------------
// Compressed 8-wide BVH node, 80 B = 5 x 16 B (layout after Ylitie, Karras, Laine 2017):
//   q0: p.x p.y p.z | ex ey ez imask
//   q1: child_base | tri_base | meta[0..3] | meta[4..7]
//   q2: qlox[0..3] qlox[4..7] qloy[0..3] qloy[4..7]
//   q3: qloz[0..3] qloz[4..7] qhix[0..3] qhix[4..7]
//   q4: qhiy[0..3] qhiy[4..7] qhiz[0..3] qhiz[4..7]
// meta: 0 = empty; internal = 0b001_xxxxx with xxxxx = 24 + slot; leaf = unary tri count (001,011,111)
// in the top 3 bits, low 5 bits = first triangle offset from tri_base (0..23).
// ---------------------------------------------------------------------------------------------
struct u4 { uint32_t x, y, z, w; };

struct RayCtx {
  vec3 o, d;       // exact origin/direction (triangle test)
  vec3 idir;       // 1/d with |d| clamped away from 0 (box tests only)
  vec3 far_bias;   // 2^100 on axes where d == 0: the slab then passes iff lo <= o (inclusive far face)
  uint32_t octinv; // 7 - octant, octant bit k set <=> d[k] < 0
  float tnear;
};

#define SPRAY_IDIR_CLAMP 8.271806125530277e-25f   // 2^-80
SPRAY_HD RayCtx make_ray_ctx(vec3 o, vec3 d, float tnear) {
  RayCtx r;
  r.o = o; r.d = d; r.tnear = tnear;
  const float eps = SPRAY_IDIR_CLAMP;
  float dx = fabsf(d.x) > eps ? d.x : copysignf(eps, d.x);
  float dy = fabsf(d.y) > eps ? d.y : copysignf(eps, d.y);
  float dz = fabsf(d.z) > eps ? d.z : copysignf(eps, d.z);
  r.idir = mk3(1.0f / dx, 1.0f / dy, 1.0f / dz);
  // A ray lying exactly in a box face (d[k] == 0, o[k] == hi[k]) would see t_far == 0 < tnear and skip
  // the box; pushing the far plane of parallel axes to +2^100 keeps such boxes (conservative).
  const float huge = 1.2676506e30f;
  r.far_bias = mk3(d.x == 0.0f ? huge : 0.0f, d.y == 0.0f ? huge : 0.0f, d.z == 0.0f ? huge : 0.0f);
  uint32_t oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
  r.octinv = 7u - oct;
  return r;
}

// 1/d exactly as domain_slab wants it: idir IS div_rn(1, d) on every axis that make_ray_ctx did not clamp
// (|d| > 2^-80); the clamped axes (zero / denormal-small components) take the division.
SPRAY_HD float exact_inv(float d, float idir) {
  return fabsf(d) > SPRAY_IDIR_CLAMP ? idir : div_rn(1.0f, d);
}

SPRAY_HD uint32_t byte_of(uint32_t w, int j) { return (w >> (8 * j)) & 0xffu; }

// Quantised plane byte j of w as the float 32768 + q, built by ONE byte permute: 0x4700qq00 has exponent
// 2^15 and q in mantissa bits 8..15 (weight 2^0).  The 32768 is folded into the FMA offset of the slab
// test (node8_hits), so a plane distance costs PRMT + FFMA and never touches the quarter-rate XU pipe
// (ncu on the first build: 48 I2F per node saturated it).
#define SPRAY_QBIAS 32768.0f
// k47 = 0x47000000 held in a REGISTER (opaque_k47) so that the selector can be the instruction's immediate:
// with both constant, ptxas makes 0x47000000 the immediate and re-materialises the selector per PRMT.
#if defined(__CUDACC__)
static __constant__ uint32_t spray_k47_const = 0x47000000u;  // constant bank: ptxas cannot fold it
#endif
SPRAY_HD uint32_t opaque_k47() {
#if defined(__CUDA_ARCH__)
  return spray_k47_const;
#else
  return 0x47000000u;
#endif
}
SPRAY_HD float byte_to_biased_float(uint32_t w, int j, uint32_t k47) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(__byte_perm(w, k47, 0x7404u | ((uint32_t)j << 4)));
#else
  return u2f(k47 | (byte_of(w, j) << 8));
#endif
}

// ---------------------------------------------------------------------------------------------
// SPRAY_POP_CULL (A/B knob, default off; CPU measurement in DESIGN.md section 6: 24-29 % fewer node visits).
// The node test also returns ONE BYTE per child slot: the child's entry distance cmin truncated to a minifloat
// (5 exponent bits, 3 mantissa bits: float bits 27..20, exponent bit 4 flipped so that bytes order like the
// distances for 2^-15 <= t < 2^17).  A (child base, mask) group keeps its two distance words, and children whose
// byte exceeds the byte of the current best hit are dropped from the mask without a node fetch: when the group
// is popped, and when the best hit improves.  Conservative by construction: truncation rounds the child's
// distance DOWN, the best hit's byte is compared with '>', out-of-range distances wrap to SMALLER bytes on the
// child side and switch the pruning off on the best-hit side.
// ---------------------------------------------------------------------------------------------
#ifndef SPRAY_POP_CULL
#define SPRAY_POP_CULL 0
#endif
#if SPRAY_POP_CULL
// per byte: x > y (unsigned); result has bit 7 of each byte set where true
SPRAY_HD uint32_t gt_u8x4(uint32_t x, uint32_t y) {
  const uint32_t d = (y | 0x80808080u) - (x & 0x7f7f7f7fu);     // bit 7 of each byte: low7(x) <= low7(y)
  return ((x & ~y) | (~(x ^ y) & ~d)) & 0x80808080u;
}
SPRAY_HD uint32_t gather_bit7(uint32_t flags) {                 // bytes 0..3 -> bits 0..3
  return ((((flags >> 7) & 0x01010101u) * 0x01020408u) >> 24) & 0xfu;
}
// drops from a group's pending-children mask (bits 31..24 of ngy, bit 24+k <-> slot k ^ octinv) every child whose
// distance byte says it starts behind best_t
SPRAY_HD uint32_t prune_group(uint32_t ngy, uint32_t d0, uint32_t d1, float best_t, uint32_t octinv) {
  if (!(best_t < 65536.0f)) return ngy;                         // INF / far hits: bytes would wrap
  const uint32_t tq = ((((f2u(best_t) >> 20) ^ 0x80u) & 0xffu)) * 0x01010101u;
  uint32_t m = gather_bit7(gt_u8x4(d0, tq)) | (gather_bit7(gt_u8x4(d1, tq)) << 4);    // stale SLOTS
  if (octinv & 1u) m = ((m & 0x55u) << 1) | ((m >> 1) & 0x55u);                        // slot k -> bit k ^ octinv
  if (octinv & 2u) m = ((m & 0x33u) << 2) | ((m >> 2) & 0x33u);
  if (octinv & 4u) m = ((m & 0x0fu) << 4) | ((m >> 4) & 0x0fu);
  return ngy & ~(m << 24);
}
#endif

// returns the 32-bit hit mask of one node: bits 31..24 internal children in traversal priority order,
// bits 23..0 one bit per leaf triangle.
//
// Slab distances: plane = p + q * 2^e  =>  t = (plane - o) * idir = q * a + b with a = 2^e * idir and
// b = (p - o) * idir.  With f = 32768 + q:  t = fma(f, a, b - 32768 * a).  The extra rounding of the
// folded offset is <= 2^-9 quantisation steps; encode_node8 keeps every child plane >= 2^-7 step inside
// its quantised plane, so the test stays conservative.  The far side carries the 1e-5 relative slack
// (FP32 rounding of the plane distances, and rays the FP32 triangle test accepts although they graze
// just outside the exact triangle) pre-multiplied into its coefficients.
// The two halves (children 0..3 / 4..7) are written as a loop that is unrolled by default (see below).
SPRAY_HD uint32_t node8_hits(const u4& q0, const u4& q1, const u4& q2, const u4& q3, const u4& q4, const RayCtx& r,
                             float tmax
#if SPRAY_POP_CULL
                             , uint32_t* dist0, uint32_t* dist1      // out: distance byte of slots 0..3 / 4..7
#endif
                             ) {
  const uint32_t e = q0.w;
  const float slack = 1.00001f;
  const float aix = u2f((e & 0xffu) << 23) * r.idir.x;
  const float aiy = u2f(((e >> 8) & 0xffu) << 23) * r.idir.y;
  const float aiz = u2f(((e >> 16) & 0xffu) << 23) * r.idir.z;
  const float ox = fmaf(-SPRAY_QBIAS, aix, (u2f(q0.x) - r.o.x) * r.idir.x);
  const float oy = fmaf(-SPRAY_QBIAS, aiy, (u2f(q0.y) - r.o.y) * r.idir.y);
  const float oz = fmaf(-SPRAY_QBIAS, aiz, (u2f(q0.z) - r.o.z) * r.idir.z);
  const float afx = aix * slack, afy = aiy * slack, afz = aiz * slack;
  const float fx = fmaf(ox, slack, r.far_bias.x), fy = fmaf(oy, slack, r.far_bias.y), fz = fmaf(oz, slack, r.far_bias.z);
  const float tfar = tmax * slack;
  const uint32_t octinv4 = r.octinv * 0x01010101u;
  const bool nx = r.idir.x < 0.0f, ny = r.idir.y < 0.0f, nz = r.idir.z < 0.0f;
  uint32_t hitmask = 0;
#if SPRAY_POP_CULL
  uint32_t dd = 0u;
#endif
  const uint32_t k47 = opaque_k47();
  // Both halves unrolled: the 12 per-half word selects fold away.  (The first fused kernel kept this a real loop —
  // ncu then showed ~40 % instruction-fetch stalls with everything unrolled; with the lean kernel variant and the
  // flat warp-synchronous loop the unrolled form wins: C2 6742 -> 7042 Mrays/s, C3 1041 -> 1087, profiles/r01x.)
#if defined(__CUDA_ARCH__) && defined(SPRAY_NODE_ROLLED)
#pragma unroll 1
#elif defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int half = 0; half < 2; ++half) {
    const uint32_t meta4 = half ? q1.w : q1.z;
    const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
    const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;
    const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
    const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
    const uint32_t lox = half ? q2.y : q2.x, loy = half ? q2.w : q2.z, loz = half ? q3.y : q3.x;
    const uint32_t hix = half ? q3.w : q3.z, hiy = half ? q4.y : q4.x, hiz = half ? q4.w : q4.z;
    const uint32_t nearx = nx ? hix : lox, farx = nx ? lox : hix;
    const uint32_t neary = ny ? hiy : loy, fary = ny ? loy : hiy;
    const uint32_t nearz = nz ? hiz : loz, farz = nz ? loz : hiz;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float t0x = fmaf(byte_to_biased_float(nearx, j, k47), aix, ox);
      const float t0y = fmaf(byte_to_biased_float(neary, j, k47), aiy, oy);
      const float t0z = fmaf(byte_to_biased_float(nearz, j, k47), aiz, oz);
      const float t1x = fmaf(byte_to_biased_float(farx, j, k47), afx, fx);
      const float t1y = fmaf(byte_to_biased_float(fary, j, k47), afy, fy);
      const float t1z = fmaf(byte_to_biased_float(farz, j, k47), afz, fz);
      const float cmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, r.tnear));
      const float cmax = fminf(fminf(t1x, t1y), fminf(t1z, tfar));
      // branch-free: a per-child branch costs a fetch redirect and a reconvergence barrier
      const uint32_t sh = byte_of(child_bits4, j) << byte_of(bit_index4, j);
      hitmask |= (cmin <= cmax) ? sh : 0u;
#if SPRAY_POP_CULL
      dd |= ((f2u(cmin) >> 20) & 0xffu) << (8 * j);
#endif
    }
#if SPRAY_POP_CULL
    // tnear below 2^-15 (a caller's own tmin) would wrap upwards: such rays keep zero bytes = never pruned
    dd = r.tnear >= 3.0517578125e-05f ? (dd ^ 0x80808080u) : 0u;
    if (half) *dist1 = dd; else *dist0 = dd;
    dd = 0u;
#endif
  }
  return hitmask;
}

// Depth bound of the per-ray traversal stack: one (child base, pending-children mask) group per BVH8 LEVEL on the
// current root-to-node path.  spraygpu_scene_commit measures the BVH8 depth (the number of collapse levels) and
// refuses trees deeper than this (SPRAYGPU_ERR_LIMIT), so a push can never run out of room at trace time; the guard
// below turns the impossible case into a trap instead of a silently dropped subtree (= missed hits).
// (30-bit Morton keys + index tie-break bound the LBVH depth by 30 + log2(n) <= 53; the 8-wide collapse consumes
// >= 1 binary level per node on the Morton part and ~3 on the balanced tie-break part: <= ~38 levels.)
#define SPRAY_STACK_SIZE 48
#if defined(__CUDA_ARCH__)
#define SPRAY_STACK_OVERFLOW() __trap()
#else
#include <stdlib.h>
#define SPRAY_STACK_OVERFLOW() abort()
#endif

struct Closest {
  float t, u, v;
  uint32_t prim; // valid iff t < inf
};

// ONE traversal routine for both query kinds (a single code instance keeps the persistent kernel's hot
// loop inside the instruction cache):
//   any == false: closest hit in ONE domain's BVH8 = lexicographic min of (t, primID) over every triangle
//                 passing S1; `best` is updated in place (seed best.t with the best t so far).
//   any == true : returns true at the first triangle passing S1 (best is untouched).
// nodes: 5 x u4 per node; tris: 3 x float4-equivalent (12 floats) per triangle.  LOADN / LOADT abstract
// the memory access (global loads on the device, plain reads in host emulation).
template <typename NodeLoader, typename TriLoader>
SPRAY_HD bool bvh8_trace(const NodeLoader& loadn, const TriLoader& loadt, const RayCtx& r, const bool any,
                         Closest& best, uint32_t* stack_x, uint32_t* stack_y, uint32_t* n_nodes,
                         uint32_t* n_tris) {
  uint32_t ngx = 0, ngy = 0x80000000u;
  int sp = 0;
#if SPRAY_POP_CULL
  uint32_t gd0 = 0u, gd1 = 0u;                                  // distance bytes of the current group
  uint32_t stack_d0[SPRAY_STACK_SIZE], stack_d1[SPRAY_STACK_SIZE];
#endif
  for (;;) {
    uint32_t tgx = 0, tgy = 0;
    if (ngy > 0x00FFFFFFu) {
      const uint32_t imask = ngy & 0xffu;
      const int bit = 31 - clz32(ngy);
      ngy &= ~(1u << bit);
      if (ngy > 0x00FFFFFFu) {
        if (sp >= SPRAY_STACK_SIZE) SPRAY_STACK_OVERFLOW();
#if SPRAY_POP_CULL
        stack_d0[sp] = gd0; stack_d1[sp] = gd1;
#endif
        stack_x[sp] = ngx; stack_y[sp] = ngy; ++sp;
      }
      const uint32_t slot = (uint32_t)(bit - 24) ^ r.octinv;
      const uint32_t rel = (uint32_t)popc32(imask & ~(0xFFFFFFFFu << slot));
      u4 q0, q1, q2, q3, q4;
      loadn(ngx + rel, q0, q1, q2, q3, q4);
      if (n_nodes) ++*n_nodes;
#if SPRAY_POP_CULL
      const uint32_t hm = node8_hits(q0, q1, q2, q3, q4, r, best.t, &gd0, &gd1);
#else
      const uint32_t hm = node8_hits(q0, q1, q2, q3, q4, r, best.t);
#endif
      ngx = q1.x; tgx = q1.y;
      ngy = (hm & 0xFF000000u) | (q0.w >> 24);
      tgy = hm & 0x00FFFFFFu;
    }
    while (tgy) {
      const int ti = ffs32(tgy) - 1;
      tgy &= tgy - 1;
      vec3 v0, e1, e2; uint32_t prim;
      loadt(tgx + (uint32_t)ti, v0, e1, e2, prim);
      if (n_tris) ++*n_tris;
      float T, U, V, ad;
      if (tri_test_raw(r.o, r.d, v0, e1, e2, r.tnear, &T, &U, &V, &ad)) {
        if (any) return true;
        const float t = div_rn(T, ad);
        if (t < best.t || (t == best.t && prim < best.prim)) {
          best.t = t; best.u = div_rn(U, ad); best.v = div_rn(V, ad); best.prim = prim;
#if SPRAY_POP_CULL
          ngy = prune_group(ngy, gd0, gd1, best.t, r.octinv);
#endif
        }
      }
    }
#if SPRAY_POP_CULL
    while (ngy <= 0x00FFFFFFu && sp > 0) {
      --sp; ngx = stack_x[sp]; gd0 = stack_d0[sp]; gd1 = stack_d1[sp];
      ngy = prune_group(stack_y[sp], gd0, gd1, best.t, r.octinv);
    }
    if (ngy <= 0x00FFFFFFu) break;
#else
    if (ngy <= 0x00FFFFFFu) {
      if (sp == 0) break;
      --sp; ngx = stack_x[sp]; ngy = stack_y[sp];
    }
#endif
  }
  return false;
}

template <typename NodeLoader, typename TriLoader>
SPRAY_HD void bvh8_closest(const NodeLoader& loadn, const TriLoader& loadt, const RayCtx& r, Closest& best,
                           uint32_t* stack_x, uint32_t* stack_y, uint32_t* n_nodes, uint32_t* n_tris) {
  bvh8_trace(loadn, loadt, r, false, best, stack_x, stack_y, n_nodes, n_tris);
}

template <typename NodeLoader, typename TriLoader>
SPRAY_HD bool bvh8_any(const NodeLoader& loadn, const TriLoader& loadt, const RayCtx& r, uint32_t* stack_x,
                       uint32_t* stack_y, uint32_t* n_nodes, uint32_t* n_tris) {
  Closest none; none.t = INFINITY; none.u = none.v = 0.f; none.prim = 0xFFFFFFFFu;
  return bvh8_trace(loadn, loadt, r, true, none, stack_x, stack_y, n_nodes, n_tris);
}

// ---------------------------------------------------------------------------------------------
// LBVH (Karras 2012) over sorted 32-bit Morton keys; ties broken by index.
// ---------------------------------------------------------------------------------------------
SPRAY_HD int lbvh_delta(const uint32_t* keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  uint32_t a = keys[i], b = keys[j];
  if (a == b) return 32 + clz32((uint32_t)i ^ (uint32_t)j);
  return clz32(a ^ b);
}

// internal node i of n-1: outputs child refs (>=0 internal index, <0: ~leaf_position) and its key range
SPRAY_HD void lbvh_node(const uint32_t* keys, int n, int i, int* left, int* right, int* first, int* last) {
  int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
  int dmin = lbvh_delta(keys, n, i, i - d);
  int lmax = 2;
  while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  int j = i + l * d;
  int dnode = lbvh_delta(keys, n, i, j);
  int s = 0;
  int t = l;
  do {
    t = (t + 1) >> 1;
    if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  int gamma = i + s * d + (d < 0 ? -1 : 0);
  int lo = i < j ? i : j, hi = i < j ? j : i;
  *left = (lo == gamma) ? ~gamma : gamma;
  *right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
  *first = lo; *last = hi;
}

// ---------------------------------------------------------------------------------------------
// Collapse one binary subtree root into one compressed 8-wide node.
// Inputs (binary LBVH, n leaves, n-1 internal nodes):
//   bl/br child refs, bfirst/blast ranges, blo/bhi internal boxes (3 floats each, stride 3),
//   llo/lhi leaf boxes by sorted position.
// Outputs: the 80-byte node, the list of internal children (binary refs, in slot order) and the leaf
//   triangle ranges (sorted positions) in slot order.
// ---------------------------------------------------------------------------------------------
#define SPRAY_LEAF_MAX 3
// CPU evaluation through the host emulation (tools/bvh_quality.py, 40 k incoherent rays): triangles tested per ray
// 5.13 -> 4.45 on wavelet.ply and 19.9 -> 12.2 on a 128 k-triangle soup, identical nodes, node visits and hit keys.
#ifndef SPRAY_SPLIT_LEAVES
#define SPRAY_SPLIT_LEAVES 1
#endif

struct CollapseOut {
  int n_inner;          // number of internal children
  int inner_ref[8];     // binary internal-node index of each, in increasing slot order
  int n_tris;           // total leaf triangles (<= 24)
  int tri_first[8];     // per leaf child (slot order): first sorted position
  int tri_count[8];     // per leaf child: 1..3
  int n_leaf;
  uint8_t meta[8];
  uint8_t imask;
  float lo[8][3], hi[8][3]; // child boxes by slot (valid where meta != 0)
};

SPRAY_HD float box_half_area(const float* lo, const float* hi) {
  float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
  return dx * dy + dy * dz + dz * dx;
}

SPRAY_HD void collapse_select(int b, const int* bl, const int* br, const int* bfirst, const int* blast,
                              const float* blo, const float* bhi, const float* llo, const float* lhi,
                              CollapseOut* out) {
  int ref[8]; float area[8]; int cnt[8];
  int nc = 0;
  {
    int c0 = bl[b], c1 = br[b];
    int cs[2] = {c0, c1};
    for (int k = 0; k < 2; ++k) {
      int c = cs[k];
      ref[nc] = c;
      if (c < 0) { cnt[nc] = 1; area[nc] = 0.f; }
      else { cnt[nc] = blast[c] - bfirst[c] + 1; area[nc] = box_half_area(blo + 3 * c, bhi + 3 * c); }
      ++nc;
    }
  }
  while (nc < 8) {
    int best = -1; float ba = -1.f;
    for (int k = 0; k < nc; ++k)
      if (ref[k] >= 0 && cnt[k] > SPRAY_LEAF_MAX && area[k] > ba) { ba = area[k]; best = k; }
    if (best < 0) break;
    int c = ref[best];
    int cs[2] = {bl[c], br[c]};
    for (int k = 0; k < 2; ++k) {
      int cc = cs[k];
      int dst = (k == 0) ? best : nc;
      ref[dst] = cc;
      if (cc < 0) { cnt[dst] = 1; area[dst] = 0.f; }
      else { cnt[dst] = blast[cc] - bfirst[cc] + 1; area[dst] = box_half_area(blo + 3 * cc, bhi + 3 * cc); }
    }
    ++nc;
  }
#if SPRAY_SPLIT_LEAVES
  // spare slots: split leaf subtrees (2..SPRAY_LEAF_MAX triangles) further, largest box first.  Same nodes, same
  // triangles, but each leaf slot's box is tighter, so fewer triangles are fetched and tested per ray; the node test
  // costs the same whether a slot is used or not.
  while (nc < 8) {
    int best = -1; float ba = -1.f;
    for (int k = 0; k < nc; ++k)
      if (ref[k] >= 0 && cnt[k] <= SPRAY_LEAF_MAX && area[k] > ba) { ba = area[k]; best = k; }
    if (best < 0) break;
    int c = ref[best];
    int cs[2] = {bl[c], br[c]};
    for (int k = 0; k < 2; ++k) {
      int cc = cs[k];
      int dst = (k == 0) ? best : nc;
      ref[dst] = cc;
      if (cc < 0) { cnt[dst] = 1; area[dst] = 0.f; }
      else { cnt[dst] = blast[cc] - bfirst[cc] + 1; area[dst] = box_half_area(blo + 3 * cc, bhi + 3 * cc); }
    }
    ++nc;
  }
#endif
  // child boxes + centroids
  float clo[8][3], chi[8][3];
  float nlo[3] = {INFINITY, INFINITY, INFINITY}, nhi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int k = 0; k < nc; ++k) {
    const float* lo; const float* hi;
    if (ref[k] < 0) { lo = llo + 3 * (~ref[k]); hi = lhi + 3 * (~ref[k]); }
    else { lo = blo + 3 * ref[k]; hi = bhi + 3 * ref[k]; }
    for (int a = 0; a < 3; ++a) {
      clo[k][a] = lo[a]; chi[k][a] = hi[a];
      nlo[a] = fminf(nlo[a], lo[a]); nhi[a] = fmaxf(nhi[a], hi[a]);
    }
  }
  // greedy octant slot assignment: maximise sum of dot(child centre - node centre, dir(slot))
  float cost[8][8];
  for (int k = 0; k < nc; ++k) {
    float cx = 0.5f * (clo[k][0] + chi[k][0]) - 0.5f * (nlo[0] + nhi[0]);
    float cy = 0.5f * (clo[k][1] + chi[k][1]) - 0.5f * (nlo[1] + nhi[1]);
    float cz = 0.5f * (clo[k][2] + chi[k][2]) - 0.5f * (nlo[2] + nhi[2]);
    for (int s = 0; s < 8; ++s)
      cost[k][s] = ((s & 1) ? cx : -cx) + ((s & 2) ? cy : -cy) + ((s & 4) ? cz : -cz);
  }
  int slot_of[8]; int child_at[8];
  for (int s = 0; s < 8; ++s) child_at[s] = -1;
  for (int k = 0; k < 8; ++k) slot_of[k] = -1;
  for (int it = 0; it < nc; ++it) {
    int bk = -1, bs = -1; float bc = -INFINITY;
    for (int k = 0; k < nc; ++k) {
      if (slot_of[k] >= 0) continue;
      for (int s = 0; s < 8; ++s) {
        if (child_at[s] >= 0) continue;
        if (cost[k][s] > bc) { bc = cost[k][s]; bk = k; bs = s; }
      }
    }
    if (bk < 0) {
      // NaN vertices make every cost NaN (no comparison succeeds): first unplaced child -> first free slot
      for (int k = 0; k < nc && bk < 0; ++k) if (slot_of[k] < 0) bk = k;
      for (int s = 0; s < 8 && bs < 0; ++s) if (child_at[s] < 0) bs = s;
    }
    slot_of[bk] = bs; child_at[bs] = bk;
  }
  out->n_inner = 0; out->n_tris = 0; out->n_leaf = 0; out->imask = 0;
  for (int s = 0; s < 8; ++s) {
    out->meta[s] = 0;
    int k = child_at[s];
    if (k < 0) continue;
    for (int a = 0; a < 3; ++a) { out->lo[s][a] = clo[k][a]; out->hi[s][a] = chi[k][a]; }
    if (ref[k] >= 0 && cnt[k] > SPRAY_LEAF_MAX) {
      out->meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
      out->imask |= (uint8_t)(1u << s);
      out->inner_ref[out->n_inner++] = ref[k];
    } else {
      int first = ref[k] < 0 ? ~ref[k] : bfirst[ref[k]];
      int c = cnt[k];
      uint32_t unary = c == 1 ? 1u : (c == 2 ? 3u : 7u);
      out->meta[s] = (uint8_t)((unary << 5) | (uint32_t)out->n_tris);
      out->tri_first[out->n_leaf] = first;
      out->tri_count[out->n_leaf] = c;
      out->n_leaf++;
      out->n_tris += c;
    }
  }
}

// Quantise child boxes and emit the 5 x u4 words.  Per axis the frame is  plane(q) = p + q * 2^e  with
// p = nlo - 2^e (one step below the node box) and 253 * 2^e >= extent, so every child plane lies in
// [plane(1), plane(254)] and can be widened.  Guarantee kept here and relied on by node8_hits:
//   plane(qlo) + 2^e/128 <= child.lo   and   plane(qhi) - 2^e/128 >= child.hi   (FP32-evaluated)
// which absorbs the <= 2^-9-step rounding of the folded slab offset and the ulp of (p - o) * idir.
// starting exponent of one axis: 2^e > ext / 253, and the step stays well above the ulp of the coordinates (or
// p = nlo - step would round away)
SPRAY_HD int quant_exp_init(float nlo, float nhi) {
  const float ext = nhi - nlo;
  int e;
  if (!(ext > 0.0f)) e = -126;
  else {
    float f = frexpf(ext / 253.0f, &e);  // ext/253 = f * 2^e, f in [0.5,1)  =>  2^e > ext/253
    (void)f;
  }
  {
    const float m = fmaxf(fabsf(nlo), fabsf(nhi));
    int em = -126;
    if (m > 0.0f) { (void)frexpf(m, &em); em -= 22; }  // 2^em >= 2 ulp(m)
    if (e < em) e = em;
  }
  if (e < -100) e = -100;
  if (e > 100) e = 100;
  return e;
}
// one child interval [clo, chi] on the grid p + q * scale: widest-safe bytes, false when the margin cannot be kept
SPRAY_HD bool quant_child(float clo, float chi, float p, float scale, uint32_t* qlo, uint32_t* qhi) {
  const float margin = 0.0078125f;  // 2^-7 step
  int li = (int)fminf(fmaxf(floorf((clo - p) / scale), 0.f), 255.f);
  int hi = (int)fminf(fmaxf(ceilf((chi - p) / scale), 0.f), 255.f);
  while (li > 0 && p + ((float)li + margin) * scale > clo) --li;
  while (hi < 255 && p + ((float)hi - margin) * scale < chi) ++hi;
  *qlo = (uint32_t)li; *qhi = (uint32_t)hi;
  return !(p + ((float)li + margin) * scale > clo || p + ((float)hi - margin) * scale < chi);
}

SPRAY_HD void encode_node8(const float* nlo, const float* nhi, const CollapseOut& c, uint32_t child_base,
                           uint32_t tri_base, u4* q) {
  uint32_t ebits[3]; float porg[3];
  uint32_t qlo[3][8], qhi[3][8];
  for (int a = 0; a < 3; ++a) {
    int e = quant_exp_init(nlo[a], nhi[a]);
    for (;; ++e) {
      const float scale = u2f((uint32_t)(e + 127) << 23);
      const float p = nlo[a] - scale;
      bool ok = true;
      for (int s = 0; s < 8 && ok; ++s) {
        if (c.meta[s] == 0) { qlo[a][s] = 255; qhi[a][s] = 0; continue; }
        ok = quant_child(c.lo[s][a], c.hi[s][a], p, scale, &qlo[a][s], &qhi[a][s]);
      }
      if (ok || e >= 120) { ebits[a] = (uint32_t)(e + 127); porg[a] = p; break; }
    }
  }
  auto pack4 = [](const uint32_t* v) -> uint32_t { return v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24); };
  uint32_t m[8];
  for (int s = 0; s < 8; ++s) m[s] = c.meta[s];
  q[0].x = f2u(porg[0]); q[0].y = f2u(porg[1]); q[0].z = f2u(porg[2]);
  q[0].w = ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | ((uint32_t)c.imask << 24);
  q[1].x = child_base; q[1].y = tri_base; q[1].z = pack4(m); q[1].w = pack4(m + 4);
  q[2].x = pack4(qlo[0]); q[2].y = pack4(qlo[0] + 4); q[2].z = pack4(qlo[1]); q[2].w = pack4(qlo[1] + 4);
  q[3].x = pack4(qlo[2]); q[3].y = pack4(qlo[2] + 4); q[3].z = pack4(qhi[0]); q[3].w = pack4(qhi[0] + 4);
  q[4].x = pack4(qhi[1]); q[4].y = pack4(qhi[1] + 4); q[4].z = pack4(qhi[2]); q[4].w = pack4(qhi[2] + 4);
}

// ---------------------------------------------------------------------------------------------
// Hit attributes — reference src/render/trimesh_buffer.cc:328-360 (colour: float interpolation then
// truncation to integer; shading normal: unnormalised barycentric mix) and utils/util.h:69-73 (unpack).
// ---------------------------------------------------------------------------------------------
SPRAY_HD uint32_t interp_color(uint32_t c0, uint32_t c1, uint32_t c2, float u, float v, float w) {
  uint32_t out = 0;
#pragma unroll
  for (int sh = 16; sh >= 0; sh -= 8) {
    float a = (float)((c0 >> sh) & 0xffu), b = (float)((c1 >> sh) & 0xffu), c = (float)((c2 >> sh) & 0xffu);
    float s = add_rn(add_rn(mul_rn(a, w), mul_rn(b, u)), mul_rn(c, v));
    out |= ((uint32_t)s) << sh;
  }
  return out;
}
SPRAY_HD float unpack_channel(uint32_t c8) { return (float)((double)c8 * 0.00392156862745098); }
SPRAY_HD vec3 unpack_color(uint32_t c) {
  return mk3(unpack_channel((c >> 16) & 0xffu), unpack_channel((c >> 8) & 0xffu), unpack_channel(c & 0xffu));
}

// ---------------------------------------------------------------------------------------------
// Sampling / BSDF pieces — reference src/render/sampler.h:49-128, sampler.cc:54-60,
// reflection.h:126-214, light.h:41-89.
// ---------------------------------------------------------------------------------------------
// The three transcendentals of the shading path (cos / sin of the disk angle, the Blinn-Phong power).  The reference
// calls glibc's cosf / sinf / powf (src/render/sampler.cc:54-60, src/render/reflection.h:202-214), which are correctly
// rounded for all practical purposes; libdevice's float versions are 1-2 ulp off, and on a multi-bounce or AO path a few
// ulp in a sample direction flips an occasional shadow / bounce decision (round 2: RMSE 1.4e-4 .. 2.7e-4 against the
// CPU checker at 384^2).  Both sides — here and the CPU restatement the tests check against — evaluate in DOUBLE and round once to float:
// the double results differ by an ulp of double at most, which changes the float in ~1 of 2^28 calls.  A few calls per
// hit; the FP64 rate is irrelevant next to the traversal.
SPRAY_HD float cos_f64(float x) { return (float)cos((double)x); }
SPRAY_HD float sin_f64(float x) { return (float)sin((double)x); }
SPRAY_HD float pow_f64(float x, float y) { return (float)pow((double)x, (double)y); }
SPRAY_HD void concentric_disk(float u1, float u2, float* dx, float* dy) {
  float r, theta;
  float sx = sub_rn(mul_rn(2.0f, u1), 1.0f);
  float sy = sub_rn(mul_rn(2.0f, u2), 1.0f);
  if (sx == 0.0f && sy == 0.0f) { *dx = 0.0f; *dy = 0.0f; return; }
  if (sx >= -sy) {
    if (sx > sy) { r = sx; theta = (sy > 0.0f) ? div_rn(sy, r) : add_rn(8.0f, div_rn(sy, r)); }
    else { r = sy; theta = sub_rn(2.0f, div_rn(sx, r)); }
  } else {
    if (sx <= sy) { r = -sx; theta = sub_rn(4.0f, div_rn(sy, r)); }
    else { r = -sy; theta = add_rn(6.0f, div_rn(sx, r)); }
  }
  theta = (float)((double)theta * (3.14159265358979323846 / 4.0)); // theta *= SPRAY_PI / 4.f (double)
  *dx = mul_rn(r, cos_f64(theta));
  *dy = mul_rn(r, sin_f64(theta));
}
SPRAY_HD vec3 local_to_world(vec3 N, vec3 v) {
  vec3 dx0 = mk3(0.f, N.z, -N.y);
  vec3 dx1 = mk3(-N.z, 0.f, N.x);
  vec3 dx = gnormalize(gdot(dx0, dx0) > gdot(dx1, dx1) ? dx0 : dx1);
  vec3 dy = gnormalize(gcross(N, dx));
  return mk3(add_rn(add_rn(mul_rn(dx.x, v.x), mul_rn(dy.x, v.y)), mul_rn(N.x, v.z)),
             add_rn(add_rn(mul_rn(dx.y, v.x), mul_rn(dy.y, v.y)), mul_rn(N.y, v.z)),
             add_rn(add_rn(mul_rn(dx.z, v.x), mul_rn(dy.z, v.y)), mul_rn(N.z, v.z)));
}
SPRAY_HD void cosine_hemisphere(float u1, float u2, vec3 N, vec3* wi, float* pdf) {
  vec3 v;
  concentric_disk(u1, u2, &v.x, &v.y);
  v.z = sqrt_rn(fmaxf(0.f, sub_rn(sub_rn(1.f, mul_rn(v.x, v.x)), mul_rn(v.y, v.y))));
  v = gnormalize(v);
  *wi = gnormalize(local_to_world(N, v));
  *pdf = mul_rn(v.z, SPRAY_ONE_OVER_PI);
}
SPRAY_HD vec3 blinn_phong(float costheta, vec3 kd, vec3 ks, float shininess, vec3 li, vec3 wi, vec3 n, vec3 wo) {
  vec3 half_hat = gnormalize(gadd(wi, wo));
  float n_dot_h = gclamp(gdot(n, half_hat), 0.0f, 1.0f);
  vec3 cs = gscale(ks, pow_f64(n_dot_h, shininess));
  vec3 cd = gscale(kd, costheta);
  return gmul(li, gadd(cd, cs));
}
SPRAY_HD bool has_positive(vec3 v) { return v.x > 0.0f || v.y > 0.0f || v.z > 0.0f; }

SPRAY_HD float fr_dielectric(float cosI, float etaI, float etaT, vec3 wo, vec3 n, vec3* wt) {
  float sin2I = fmaxf(0.0f, sub_rn(1.0f, mul_rn(cosI, cosI)));
  float eta = div_rn(etaI, etaT);
  float sin2T = mul_rn(mul_rn(eta, eta), sin2I);
  if (sin2T >= 1.0f) return 1.0f;
  float cosT = sqrt_rn(sub_rn(1.0f, sin2T));
  float Rparl = div_rn(sub_rn(mul_rn(etaT, cosI), mul_rn(etaI, cosT)), add_rn(mul_rn(etaT, cosI), mul_rn(etaI, cosT)));
  float Rperp = div_rn(sub_rn(mul_rn(etaI, cosI), mul_rn(etaT, cosT)), add_rn(mul_rn(etaI, cosI), mul_rn(etaT, cosT)));
  float fr = div_rn(add_rn(mul_rn(Rparl, Rparl), mul_rn(Rperp, Rperp)), 2.0f);
  *wt = gadd(gscale(gneg(wo), eta), gscale(n, sub_rn(mul_rn(eta, cosI), cosT)));
  return fr;
}
SPRAY_HD bool refract_dir(float cosI, float etaI, float etaT, vec3 wo, vec3 n, vec3* wt) {
  float sin2I = fmaxf(0.0f, sub_rn(1.0f, mul_rn(cosI, cosI)));
  float eta = div_rn(etaI, etaT);
  float sin2T = mul_rn(mul_rn(eta, eta), sin2I);
  if (sin2T >= 1.0f) return false;
  float cosT = sqrt_rn(sub_rn(1.0f, sin2T));
  *wt = gadd(gscale(gneg(wo), eta), gscale(n, sub_rn(mul_rn(eta, cosI), cosT)));
  return true;
}

// Eye ray direction — reference src/render/camera.h:171-209
struct CameraDev {
  float pos[3], lowerleft[3], wvec[3], hvec[3];
  int image_w, image_h;
};
SPRAY_HD vec3 camera_dir(const CameraDev& c, float x, float y) {
  float u = div_rn(x, (float)c.image_w);
  float v = div_rn(y, (float)c.image_h);
  vec3 ll = mk3(c.lowerleft[0], c.lowerleft[1], c.lowerleft[2]);
  vec3 wv = mk3(c.wvec[0], c.wvec[1], c.wvec[2]);
  vec3 hv = mk3(c.hvec[0], c.hvec[1], c.hvec[2]);
  vec3 pos = mk3(c.pos[0], c.pos[1], c.pos[2]);
  return gnormalize(gsub(gadd(gadd(ll, gscale(wv, u)), gscale(hv, v)), pos));
}

} // namespace spraygpu
