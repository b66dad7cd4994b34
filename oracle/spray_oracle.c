/*
 * spray_oracle.c — CPU oracle (TEST INFRASTRUCTURE ONLY; see spray_oracle.h).  parity unpinned.
 *
 * Plain C99 + OpenMP.  Build: see oracle/Makefile (-ffp-contract=off so that ONLY the explicit fmaf()
 * calls below are fused; every other FP32 operation rounds once, in the order written).
 *
 * File:line citations are relative to /root/reference.
 */
#include "spray_oracle.h"

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------------------------------
 * vec3 helpers with glm 0.9.9.3 operation order (deps/glm/glm/detail/func_geometric.inl:48-90,
 * func_exponential.inl:134-139): dot = (x*x' + y*y') + z*z' ; normalize = v * (1/sqrt(dot(v,v))).
 * ---------------------------------------------------------------------------------------------- */
typedef struct { float x, y, z; } v3;
static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vmul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 vscale(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline float gdot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline v3 gcross(v3 x, v3 y) {
  return V(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
static inline v3 gnormalize(v3 v) { return vscale(v, 1.0f / sqrtf(gdot(v, v))); }
static inline float gclamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
static inline float glength(v3 v) { return sqrtf(gdot(v, v)); }

/* Embree-style fused helpers used ONLY by the triangle test (FP32 contract S1). */
static inline float edot(v3 a, v3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
static inline v3 ecross(v3 a, v3 b) {
  return V(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

#define RAY_EPS 0.001f /* SPRAY_RAY_EPSILON, render/spray.h:46 */
#define ONE_OVER_PI 0.3183098861837907f /* render/spray.h:50 */

/* ------------------------------------------------------------------------------------------------
 * RNG — deps/embree/random_sampler.h:34-108
 * ---------------------------------------------------------------------------------------------- */
static inline uint32_t murmur_mix(uint32_t hash, uint32_t k) {
  k *= 0xcc9e2d51u;
  k = (k << 15) | (k >> 17);
  k *= 0x1b873593u;
  hash ^= k;
  hash = ((hash << 13) | (hash >> 19)) * 5u + 0xe6546b64u;
  return hash;
}
static inline uint32_t murmur_fin(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
uint32_t orc_rng_init1(int id) { return murmur_fin(murmur_mix(0u, (uint32_t)id)); }
uint32_t orc_rng_init2(int pixel_id, int sample_id) {
  return murmur_fin(murmur_mix(murmur_mix(0u, (uint32_t)pixel_id), (uint32_t)sample_id));
}
float orc_rng_get_float(uint32_t* s) {
  *s = *s * 1664525u + 1013904223u;            /* LCG_next, :63-68 */
  int v = (int)(*s >> 1);                      /* getInt, :92-95 */
  return (float)v * 4.656612873077392578125e-10f; /* getFloat, :102-104 */
}

/* ------------------------------------------------------------------------------------------------
 * Morton — render/morton.h:33-52
 * ---------------------------------------------------------------------------------------------- */
static inline uint32_t expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
uint32_t orc_morton(float x, float y, float z) {
  x = fminf(fmaxf(x * 1024.0f, 0.0f), 1023.0f);
  y = fminf(fmaxf(y * 1024.0f, 0.0f), 1023.0f);
  z = fminf(fmaxf(z * 1024.0f, 0.0f), 1023.0f);
  return expand_bits((uint32_t)x) * 4 + expand_bits((uint32_t)y) * 2 + expand_bits((uint32_t)z);
}

/* ------------------------------------------------------------------------------------------------
 * Slab test — render/aabb.h:139-169 verbatim in behaviour (incl. NaN / early-out order; tmin is NOT
 * clamped to t0; outputs are written only when the function does not early-out).
 * ---------------------------------------------------------------------------------------------- */
int orc_intersect_aabb(const float b[6], const float org[3], const float dir[3], float t0, float t1,
                       float* tmin_o, float* tmax_o) {
  float inv[3] = {1.0f / dir[0], 1.0f / dir[1], 1.0f / dir[2]};
  int sx = inv[0] < 0.0f, sy = inv[1] < 0.0f, sz = inv[2] < 0.0f;
  float tmin = (b[sx * 3 + 0] - org[0]) * inv[0];
  float tmax = (b[(1 - sx) * 3 + 0] - org[0]) * inv[0];
  float tymin = (b[sy * 3 + 1] - org[1]) * inv[1];
  float tymax = (b[(1 - sy) * 3 + 1] - org[1]) * inv[1];
  if ((tmin > tymax) || (tymin > tmax)) return 0;
  if (tymin > tmin) tmin = tymin;
  if (tymax < tmax) tmax = tymax;
  float tzmin = (b[sz * 3 + 2] - org[2]) * inv[2];
  float tzmax = (b[(1 - sz) * 3 + 2] - org[2]) * inv[2];
  if ((tmin > tzmax) || (tzmin > tmax)) return 0;
  if (tzmin > tmin) tmin = tzmin;
  if (tzmax < tmax) tmax = tzmax;
  *tmin_o = tmin;
  *tmax_o = tmax;
  return (tmin < t1) && (tmax > t0);
}

/* ------------------------------------------------------------------------------------------------
 * Colour: utils/util.h:69-98 (unpack = uint * double(1/255) narrowed), trimesh_buffer.cc:336-350
 * (float interpolation, truncation to uint32 before repack).
 * ---------------------------------------------------------------------------------------------- */
void orc_unpack_color(uint32_t c, float rgb[3]) {
  rgb[0] = (float)(((c >> 16) & 0xff) * 0.00392156862745098);
  rgb[1] = (float)(((c >> 8) & 0xff) * 0.00392156862745098);
  rgb[2] = (float)((c & 0xff) * 0.00392156862745098);
}
uint32_t orc_interp_color(uint32_t c0, uint32_t c1, uint32_t c2, float u, float v) {
  uint32_t rgb[9] = {(c0 >> 16) & 0xff, (c0 >> 8) & 0xff, c0 & 0xff, (c1 >> 16) & 0xff, (c1 >> 8) & 0xff,
                     c1 & 0xff,         (c2 >> 16) & 0xff, (c2 >> 8) & 0xff, c2 & 0xff};
  float w = 1.f - u - v;
  /* uint32 * float -> float ; sum left to right ; truncating conversion */
  uint32_t r = (uint32_t)(((float)rgb[0] * w) + ((float)rgb[3] * u) + ((float)rgb[6] * v));
  uint32_t g = (uint32_t)(((float)rgb[1] * w) + ((float)rgb[4] * u) + ((float)rgb[7] * v));
  uint32_t b = (uint32_t)(((float)rgb[2] * w) + ((float)rgb[5] * u) + ((float)rgb[8] * v));
  return (r << 16) | (g << 8) | b;
}

/* ------------------------------------------------------------------------------------------------
 * Camera — render/camera.h:128-209; default pose — render/spray_renderer.inl:87-105
 * ---------------------------------------------------------------------------------------------- */
void orc_camera_init(const float campos[3], const float lookat[3], const float up[3], float vfov, int image_w,
                     int image_h, OrcCamera* out) {
  float aspect = (float)image_w / (float)image_h;
  float theta = (float)(vfov * M_PI / 180.0f); /* float*double/float -> double, stored to float */
  float half_h = (float)tan(theta / 2.0f);     /* ::tan(double) on a float argument */
  float half_w = aspect * half_h;
  v3 pos = V(campos[0], campos[1], campos[2]);
  v3 w = vsub(pos, V(lookat[0], lookat[1], lookat[2]));
  v3 u = gcross(V(up[0], up[1], up[2]), w);
  v3 v = gcross(w, u);
  w = gnormalize(w); u = gnormalize(u); v = gnormalize(v);
  v3 center = vsub(pos, w);
  v3 ll = vsub(vsub(center, vscale(u, half_w)), vscale(v, half_h));
  v3 wv = vscale(u, 2.0f * half_w);
  v3 hv = vscale(v, 2.0f * half_h);
  out->pos[0] = pos.x; out->pos[1] = pos.y; out->pos[2] = pos.z;
  out->lowerleft[0] = ll.x; out->lowerleft[1] = ll.y; out->lowerleft[2] = ll.z;
  out->wvec[0] = wv.x; out->wvec[1] = wv.y; out->wvec[2] = wv.z;
  out->hvec[0] = hv.x; out->hvec[1] = hv.y; out->hvec[2] = hv.z;
  out->image_w = image_w; out->image_h = image_h;
}

void orc_default_camera_pose(const float sb[6], float campos[3], float lookat[3], float up[3]) {
  v3 extent = V(sb[3] - sb[0], sb[4] - sb[1], sb[5] - sb[2]);
  /* Aabb::getCenter, aabb.cc:133-137: 0.5 (double) * float sum */
  lookat[0] = (float)(0.5 * (sb[0] + sb[3]));
  lookat[1] = (float)(0.5 * (sb[1] + sb[4]));
  lookat[2] = (float)(0.5 * (sb[2] + sb[5]));
  float s = glength(extent) * 0.5f;
  campos[0] = lookat[0] + s * 0.0f;
  campos[1] = lookat[1] + s * 0.0f;
  campos[2] = lookat[2] + s * 1.0f;
  up[0] = 0.f; up[1] = 1.f; up[2] = 0.f;
}

static inline v3 cam_dir(const OrcCamera* c, float x, float y) {
  float u = x / (float)c->image_w; /* image_w_ is stored as float, camera.h:158 */
  float v = y / (float)c->image_h;
  v3 ll = V(c->lowerleft[0], c->lowerleft[1], c->lowerleft[2]);
  v3 wv = V(c->wvec[0], c->wvec[1], c->wvec[2]);
  v3 hv = V(c->hvec[0], c->hvec[1], c->hvec[2]);
  v3 pos = V(c->pos[0], c->pos[1], c->pos[2]);
  v3 d = vsub(vadd(vadd(ll, vscale(wv, u)), vscale(hv, v)), pos); /* camera.h:171-175 */
  return gnormalize(d);
}
void orc_camera_ray(const OrcCamera* c, float x, float y, float org[3], float dir[3]) {
  v3 d = cam_dir(c, x, y);
  org[0] = c->pos[0]; org[1] = c->pos[1]; org[2] = c->pos[2];
  dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}

/* ------------------------------------------------------------------------------------------------
 * Blocking tiles of the speculative OOC tracer — render/tile.cc:204-235 (makeVerticalStripe),
 * :317-391 (ImageScheduleTileList::init).  samid of a sample = its index inside its blocking tile.
 * ---------------------------------------------------------------------------------------------- */
int orc_image_schedule_tiles(int64_t image_w, int64_t image_h, int64_t spp, int64_t num_ranks, int rank,
                             int64_t max_samples, OrcTile* out, int cap) {
  OrcTile vs;
  int w = (int)(image_w / num_ranks); if (w < 1) w = 1;
  vs.x = rank * w;
  int xend = (int)image_w;
  if (vs.x >= xend) return 0;
  vs.w = ((vs.x + w > xend) || (rank == num_ranks - 1)) ? xend - vs.x : w;
  vs.y = 0; vs.h = (int)image_h;
  int64_t num_samples = (int64_t)vs.w * (int64_t)vs.h * spp;
  int64_t est = (num_samples + max_samples - 1) / max_samples;
  int ntiles = (int)est;
  int tile_h = vs.h / ntiles;
  if (tile_h < 1) tile_h = 1; /* reference divides by zero here; we clamp (documented deviation) */
  ntiles = (vs.h + tile_h - 1) / tile_h;
  int i = 0;
  for (int y = 0; y < vs.h; y += tile_h) {
    int h = tile_h < vs.h - y ? tile_h : vs.h - y;
    if (i < cap) { out[i].x = vs.x; out[i].y = y; out[i].w = vs.w; out[i].h = h; }
    ++i;
  }
  return i;
}

/* ------------------------------------------------------------------------------------------------
 * In-situ partition — render/data_partition.h:59-137 (GROUP_CLOSE_DOMAINS branch)
 * ---------------------------------------------------------------------------------------------- */
typedef struct { uint32_t code; int domain; } MortonCode;
static int cmp_code(const void* a, const void* b) {
  const MortonCode* x = (const MortonCode*)a; const MortonCode* y = (const MortonCode*)b;
  if (x->code != y->code) return x->code < y->code ? -1 : 1;
  return x->domain - y->domain; /* reference: unstable std::sort; ties made deterministic here */
}
void orc_partition(int nd, const float* wb, const float sb[6], int num_ranks, int* domain_to_rank,
                   uint32_t* codes_out) {
  v3 diag = V(sb[3] - sb[0], sb[4] - sb[1], sb[5] - sb[2]);
  v3 sf = V(1.0f / diag.x, 1.0f / diag.y, 1.0f / diag.z);
  v3 mn = vmul(V(sb[0], sb[1], sb[2]), sf);
  v3 off = vsub(V(0, 0, 0), mn);
  MortonCode* codes = (MortonCode*)malloc(sizeof(MortonCode) * nd);
  for (int i = 0; i < nd; ++i) {
    const float* b = wb + 6 * i;
    v3 c = V((float)(0.5 * (b[0] + b[3])), (float)(0.5 * (b[1] + b[4])), (float)(0.5 * (b[2] + b[5])));
    /* trans*scale applied to (c,1): per row (s*c + 0*..) + (0*.. + off*1) = s*c + off */
    v3 p = V(sf.x * c.x + off.x, sf.y * c.y + off.y, sf.z * c.z + off.z);
    codes[i].code = orc_morton(p.x, p.y, p.z);
    codes[i].domain = i;
    if (codes_out) codes_out[i] = codes[i].code;
  }
  qsort(codes, nd, sizeof(MortonCode), cmp_code);
  int shares = nd / num_ranks; if (shares < 1) shares = 1;
  int rank = 0, s = 0;
  for (int i = 0; i < nd; ++i) {
    domain_to_rank[codes[i].domain] = rank;
    if (++s == shares) { s = 0; if (++rank == num_ranks) rank = 0; }
  }
  free(codes);
}

/* ------------------------------------------------------------------------------------------------
 * Sampling — render/sampler.h:49-128, sampler.cc:54-60
 * ---------------------------------------------------------------------------------------------- */
static void concentric_disk(float u1, float u2, float* dx, float* dy) {
  float r, theta;
  float sx = 2 * u1 - 1;
  float sy = 2 * u2 - 1;
  if (sx == 0.0 && sy == 0.0) { *dx = 0.0f; *dy = 0.0f; return; }
  if (sx >= -sy) {
    if (sx > sy) { r = sx; if (sy > 0.0) theta = sy / r; else theta = 8.0f + sy / r; }
    else { r = sy; theta = 2.0f - sx / r; }
  } else {
    if (sx <= sy) { r = -sx; theta = 4.0f - sy / r; }
    else { r = -sy; theta = 6.0f + sx / r; }
  }
  theta = (float)(theta * (M_PI / 4.f)); /* theta *= SPRAY_PI / 4.f  (double product, narrowed) */
  /* cos / sin / pow: evaluated in double and rounded once, on BOTH sides (core.cuh cos_f64 / sin_f64 / pow_f64): equal to
   * glibc's cosf / sinf / powf (what the reference calls) except where those are not correctly rounded */
  *dx = r * (float)cos((double)theta);
  *dy = r * (float)sin((double)theta);
}
static v3 local_to_world(v3 N, v3 v) {
  v3 dx0 = V(0, N.z, -N.y);
  v3 dx1 = V(-N.z, 0, N.x);
  v3 dx = gnormalize(gdot(dx0, dx0) > gdot(dx1, dx1) ? dx0 : dx1);
  v3 dy = gnormalize(gcross(N, dx));
  /* mat3(dx,dy,N) * v : m[0][i]*v.x + m[1][i]*v.y + m[2][i]*v.z (type_mat3x3.inl:468-474) */
  return V(dx.x * v.x + dy.x * v.y + N.x * v.z, dx.y * v.x + dy.y * v.y + N.y * v.z,
           dx.z * v.x + dy.z * v.y + N.z * v.z);
}
static void cosine_hemisphere(float u1, float u2, v3 N, v3* wi, float* pdf) {
  v3 v;
  concentric_disk(u1, u2, &v.x, &v.y);
  v.z = sqrtf(fmaxf(0.f, 1.f - v.x * v.x - v.y * v.y));
  v = gnormalize(v);
  *wi = gnormalize(local_to_world(N, v));
  *pdf = v.z * ONE_OVER_PI;
}
void orc_cosine_hemisphere(float u1, float u2, const float n[3], float wi[3], float* pdf) {
  v3 w; cosine_hemisphere(u1, u2, V(n[0], n[1], n[2]), &w, pdf);
  wi[0] = w.x; wi[1] = w.y; wi[2] = w.z;
}

/* blinnPhong — render/reflection.h:202-214 */
static v3 blinn_phong(float costheta, v3 kd, v3 ks, float shininess, v3 li, v3 wi, v3 n, v3 wo) {
  v3 half_hat = gnormalize(vadd(wi, wo));
  float n_dot_h = gclamp(gdot(n, half_hat), 0.0f, 1.0f);
  v3 cs = vscale(ks, (float)pow((double)n_dot_h, (double)shininess));
  v3 cd = vscale(kd, costheta);
  return vmul(li, vadd(cd, cs));
}
void orc_blinn_phong(float costheta, const float kd[3], const float ks[3], float shininess, const float li[3],
                     const float wi[3], const float n[3], const float wo[3], float out[3]) {
  v3 r = blinn_phong(costheta, V(kd[0], kd[1], kd[2]), V(ks[0], ks[1], ks[2]), shininess, V(li[0], li[1], li[2]),
                     V(wi[0], wi[1], wi[2]), V(n[0], n[1], n[2]), V(wo[0], wo[1], wo[2]));
  out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

/* PPM quantisation — display/image.h:183-204 */
void orc_ppm_quantize(const float* rgba, int n, uint32_t* out) {
  for (int i = 0; i < n; ++i)
    for (int c = 0; c < 3; ++c)
      out[3 * i + c] = (unsigned)gclamp(rgba[4 * i + c] * 1023.f, 0.0f, 1023.0f);
}

/* LRU — render/lru_cache.cc:65-171 (list order = recency, evict front, reuse its block id) */
void orc_lru_sequence(int nd, int cap, const int* req, int n, int* hits, int* blocks) {
  int* order = (int*)malloc(sizeof(int) * cap); /* domains, LRU first */
  int* blk_of = (int*)malloc(sizeof(int) * nd);
  int size = 0;
  for (int i = 0; i < nd; ++i) blk_of[i] = -1;
  for (int k = 0; k < n; ++k) {
    int d = req[k];
    if (blk_of[d] >= 0) {
      hits[k] = 1; blocks[k] = blk_of[d];
      int p = 0; while (order[p] != d) ++p;
      memmove(order + p, order + p + 1, sizeof(int) * (size - p - 1));
      order[size - 1] = d;
    } else {
      hits[k] = 0;
      if (size < cap) { blk_of[d] = size; order[size++] = d; }
      else {
        int old = order[0];
        memmove(order, order + 1, sizeof(int) * (size - 1));
        blk_of[d] = blk_of[old]; blk_of[old] = -1;
        order[size - 1] = d;
      }
      blocks[k] = blk_of[d];
    }
  }
  free(order); free(blk_of);
}

/* ------------------------------------------------------------------------------------------------
 * FP32 contract S1 — ray/triangle test.  Structure follows the Moeller-Trumbore intersector of
 * Embree's triangle4v ([from memory of the un-vendored dependency, unverified]): no back-face culling,
 * inclusive edge tests, u,v relative to v1,v2, tnear < t (tfar = inf in every SpRay call:
 * render/rays.h:345-363,389-423).  Division is IEEE (Embree uses rcp+Newton; not reproducible).
 * ---------------------------------------------------------------------------------------------- */
static inline int tri_test(v3 O, v3 D, v3 v0, v3 v1, v3 v2, float tnear, float* t, float* u, float* v) {
  v3 e1 = vsub(v0, v1);
  v3 e2 = vsub(v2, v0);
  v3 Ng = ecross(e2, e1);
  v3 C = vsub(v0, O);
  v3 R = ecross(C, D);
  float den = edot(Ng, D);
  float absDen = fabsf(den);
  uint32_t sgn = f2u(den) & 0x80000000u;
  float U = u2f(f2u(edot(R, e2)) ^ sgn);
  float Vv = u2f(f2u(edot(R, e1)) ^ sgn);
  if (!(den != 0.0f)) return 0;
  if (!(U >= 0.0f)) return 0;
  if (!(Vv >= 0.0f)) return 0;
  if (!(U + Vv <= absDen)) return 0;
  float T = u2f(f2u(edot(Ng, C)) ^ sgn);
  if (!(T > absDen * tnear)) return 0;
  *t = T / absDen;
  *u = U / absDen;
  *v = Vv / absDen;
  return 1;
}
void orc_tri_test(const float org[3], const float dir[3], const float v0[3], const float v1[3], const float v2[3],
                  float tnear, int* hit, float* t, float* u, float* v) {
  *hit = tri_test(V(org[0], org[1], org[2]), V(dir[0], dir[1], dir[2]), V(v0[0], v0[1], v0[2]),
                  V(v1[0], v1[1], v1[2]), V(v2[0], v2[1], v2[2]), tnear, t, u, v);
}

/* double-precision truth for O1 (same algebra, plain arithmetic) */
typedef struct { double x, y, z; } d3;
static inline d3 D3(v3 a) { d3 r = {a.x, a.y, a.z}; return r; }
static inline d3 dsub(d3 a, d3 b) { d3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static inline double ddot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline d3 dcross(d3 a, d3 b) {
  d3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
  return r;
}
static inline int tri_test_f64(d3 O, d3 Dd, d3 v0, d3 v1, d3 v2, double tnear, double* t, double* u, double* v) {
  d3 e1 = dsub(v0, v1), e2 = dsub(v2, v0);
  d3 Ng = dcross(e2, e1);
  d3 C = dsub(v0, O);
  d3 R = dcross(C, Dd);
  double den = ddot(Ng, Dd);
  if (den == 0.0) return 0;
  double s = den < 0 ? -1.0 : 1.0, a = fabs(den);
  double U = s * ddot(R, e2), Vv = s * ddot(R, e1);
  if (U < 0.0 || Vv < 0.0 || U + Vv > a) return 0;
  double T = s * ddot(Ng, C);
  if (!(T > a * tnear)) return 0;
  *t = T / a; *u = U / a; *v = Vv / a;
  return 1;
}

/* ------------------------------------------------------------------------------------------------
 * Scene storage
 * ---------------------------------------------------------------------------------------------- */
typedef struct { float lo[3], hi[3]; int left; int count; } BNode; /* count>0: leaf, left=first prim slot */

typedef struct {
  size_t nv, nf;
  float* verts;    /* world space, nv*3 */
  uint32_t* faces; /* nf*3 */
  uint32_t* rgb;   /* nv */
  float* normals;  /* nv*3 */
  float bounds[6]; /* world AABB as given by the scene file (scene_loader.cc:352-357) */
  int mtl_type;
  float mtl[3];
  BNode* nodes; int nnodes;
  uint32_t* prim; /* BVH leaf order -> primID */
} Domain;

struct OrcScene { int nd; Domain* d; };

OrcScene* orc_scene_create(int nd) {
  OrcScene* s = (OrcScene*)calloc(1, sizeof(OrcScene));
  s->nd = nd; s->d = (Domain*)calloc(nd, sizeof(Domain));
  return s;
}
static void domain_free(Domain* d) {
  free(d->verts); free(d->faces); free(d->rgb); free(d->normals); free(d->nodes); free(d->prim);
  memset(d, 0, sizeof(*d));
}
void orc_scene_destroy(OrcScene* s) {
  if (!s) return;
  for (int i = 0; i < s->nd; ++i) domain_free(&s->d[i]);
  free(s->d); free(s);
}

/* computeNormals — render/trimesh_buffer.cc:267-326 (face order = summation order, no FMA) */
static void compute_normals(Domain* d) {
  float* n = d->normals;
  memset(n, 0, sizeof(float) * 3 * d->nv);
  for (size_t i = 0; i < d->nf; ++i) {
    size_t a = (size_t)d->faces[3 * i] * 3, b = (size_t)d->faces[3 * i + 1] * 3, c = (size_t)d->faces[3 * i + 2] * 3;
    v3 v0 = V(d->verts[a], d->verts[a + 1], d->verts[a + 2]);
    v3 v1 = V(d->verts[b], d->verts[b + 1], d->verts[b + 2]);
    v3 v2 = V(d->verts[c], d->verts[c + 1], d->verts[c + 2]);
    v3 u = vsub(v1, v0), v = vsub(v2, v0);
    v3 nn = V(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
    n[a] += nn.x; n[a + 1] += nn.y; n[a + 2] += nn.z;
    n[b] += nn.x; n[b + 1] += nn.y; n[b + 2] += nn.z;
    n[c] += nn.x; n[c + 1] += nn.y; n[c + 2] += nn.z;
  }
}

void orc_scene_set_domain(OrcScene* s, int id, const float* verts, size_t nv, const uint32_t* faces, size_t nf,
                          const uint32_t* rgb, const float* m, const float wb[6], int mtl_type,
                          const float mtl[3]) {
  Domain* d = &s->d[id];
  domain_free(d);
  d->nv = nv; d->nf = nf;
  d->verts = (float*)malloc(sizeof(float) * 3 * nv);
  d->faces = (uint32_t*)malloc(sizeof(uint32_t) * 3 * nf);
  d->rgb = (uint32_t*)calloc(nv, sizeof(uint32_t));
  d->normals = (float*)malloc(sizeof(float) * 3 * nv);
  memcpy(d->faces, faces, sizeof(uint32_t) * 3 * nf);
  if (rgb) memcpy(d->rgb, rgb, sizeof(uint32_t) * nv);
  for (size_t i = 0; i < nv; ++i) {
    float x = verts[3 * i], y = verts[3 * i + 1], z = verts[3 * i + 2];
    if (m) {
      /* glm mat4*vec4 (type_mat4x4.inl:536-): (m0*x + m1*y) + (m2*z + m3*1), column-major m[col*4+row] */
      d->verts[3 * i + 0] = (m[0] * x + m[4] * y) + (m[8] * z + m[12] * 1.0f);
      d->verts[3 * i + 1] = (m[1] * x + m[5] * y) + (m[9] * z + m[13] * 1.0f);
      d->verts[3 * i + 2] = (m[2] * x + m[6] * y) + (m[10] * z + m[14] * 1.0f);
    } else {
      d->verts[3 * i] = x; d->verts[3 * i + 1] = y; d->verts[3 * i + 2] = z;
    }
  }
  memcpy(d->bounds, wb, sizeof(float) * 6);
  d->mtl_type = mtl_type;
  d->mtl[0] = mtl ? mtl[0] : 1.f; d->mtl[1] = mtl ? mtl[1] : 1.f; d->mtl[2] = mtl ? mtl[2] : 1.f;
}

/* ---- binned-SAH BVH2 (oracle's own acceleration structure; any conservative BVH gives the same
 *      answers because the result is defined as the lexicographic min over ALL passing triangles) ---- */
typedef struct { float lo[3], hi[3], c[3]; } PrimBox;
typedef struct { Domain* d; PrimBox* pb; int nalloc; } Builder;

static void box_init(float lo[3], float hi[3]) {
  for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
}
static void box_grow(float lo[3], float hi[3], const float l2[3], const float h2[3]) {
  for (int k = 0; k < 3; ++k) { if (l2[k] < lo[k]) lo[k] = l2[k]; if (h2[k] > hi[k]) hi[k] = h2[k]; }
}
static float box_area(const float lo[3], const float hi[3]) {
  float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
  if (dx < 0) return 0.f;
  return dx * dy + dy * dz + dz * dx;
}
#define NBINS 16
#define LEAF_MAX 4
static int build_rec(Builder* B, int first, int count) {
  Domain* d = B->d;
  int ni = d->nnodes++;
  BNode* n = &d->nodes[ni];
  float clo[3], chi[3];
  box_init(n->lo, n->hi); box_init(clo, chi);
  for (int i = first; i < first + count; ++i) {
    PrimBox* p = &B->pb[d->prim[i]];
    box_grow(n->lo, n->hi, p->lo, p->hi);
    box_grow(clo, chi, p->c, p->c);
  }
  if (count <= LEAF_MAX) { n->left = first; n->count = count; return ni; }
  int best_axis = -1, best_bin = -1; float best_cost = FLT_MAX;
  for (int ax = 0; ax < 3; ++ax) {
    float ext = chi[ax] - clo[ax];
    if (!(ext > 0.f)) continue;
    float blo[NBINS][3], bhi[NBINS][3]; int bc[NBINS];
    for (int b = 0; b < NBINS; ++b) { box_init(blo[b], bhi[b]); bc[b] = 0; }
    float sc = (float)NBINS / ext;
    for (int i = first; i < first + count; ++i) {
      PrimBox* p = &B->pb[d->prim[i]];
      int b = (int)((p->c[ax] - clo[ax]) * sc); if (b >= NBINS) b = NBINS - 1; if (b < 0) b = 0;
      box_grow(blo[b], bhi[b], p->lo, p->hi); bc[b]++;
    }
    float ra[NBINS]; int rc[NBINS];
    float lo[3], hi[3]; box_init(lo, hi); int c = 0;
    for (int b = NBINS - 1; b > 0; --b) { box_grow(lo, hi, blo[b], bhi[b]); c += bc[b]; ra[b] = box_area(lo, hi); rc[b] = c; }
    box_init(lo, hi); c = 0;
    for (int b = 0; b < NBINS - 1; ++b) {
      box_grow(lo, hi, blo[b], bhi[b]); c += bc[b];
      if (c == 0 || rc[b + 1] == 0) continue;
      float cost = box_area(lo, hi) * c + ra[b + 1] * rc[b + 1];
      if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = b; }
    }
  }
  int mid;
  if (best_axis < 0) {
    mid = first + count / 2;
  } else {
    float ext = chi[best_axis] - clo[best_axis];
    float sc = (float)NBINS / ext;
    int i = first, j = first + count - 1;
    while (i <= j) {
      PrimBox* p = &B->pb[d->prim[i]];
      int b = (int)((p->c[best_axis] - clo[best_axis]) * sc); if (b >= NBINS) b = NBINS - 1; if (b < 0) b = 0;
      if (b <= best_bin) ++i; else { uint32_t t = d->prim[i]; d->prim[i] = d->prim[j]; d->prim[j] = t; --j; }
    }
    mid = i;
    if (mid == first || mid == first + count) mid = first + count / 2;
  }
  int l = build_rec(B, first, mid - first);
  int r = build_rec(B, mid, first + count - mid);
  n = &d->nodes[ni]; /* re-fetch: nothing moved (array is preallocated), kept for clarity */
  n->left = l;
  n->count = -r; /* internal node: count < 0, right child index = -count (r >= 2 always) */
  return ni;
}

static void build_domain(Domain* d) {
  compute_normals(d);
  free(d->nodes); free(d->prim);
  d->nodes = (BNode*)malloc(sizeof(BNode) * (2 * d->nf + 1));
  d->prim = (uint32_t*)malloc(sizeof(uint32_t) * (d->nf ? d->nf : 1));
  d->nnodes = 0;
  if (d->nf == 0) return;
  Builder B; B.d = d; B.pb = (PrimBox*)malloc(sizeof(PrimBox) * d->nf);
  for (size_t i = 0; i < d->nf; ++i) {
    PrimBox* p = &B.pb[i];
    box_init(p->lo, p->hi);
    for (int k = 0; k < 3; ++k) {
      const float* v = &d->verts[(size_t)d->faces[3 * i + k] * 3];
      box_grow(p->lo, p->hi, v, v);
    }
    for (int k = 0; k < 3; ++k) p->c[k] = 0.5f * (p->lo[k] + p->hi[k]);
    d->prim[i] = (uint32_t)i;
  }
  build_rec(&B, 0, (int)d->nf);
  free(B.pb);
}

void orc_scene_commit(OrcScene* s) {
#pragma omp parallel for schedule(dynamic, 1)
  for (int i = 0; i < s->nd; ++i) build_domain(&s->d[i]);
}
void orc_scene_bounds(const OrcScene* s, float out[6]) {
  /* Scene::mergeDomainBounds, render/scene.inl:296-349 */
  box_init(out, out + 3);
  for (int i = 0; i < s->nd; ++i) box_grow(out, out + 3, s->d[i].bounds, s->d[i].bounds + 3);
}
size_t orc_scene_num_verts(const OrcScene* s, int id) { return s->d[id].nv; }
void orc_scene_get_normals(const OrcScene* s, int id, float* out) {
  memcpy(out, s->d[id].normals, sizeof(float) * 3 * s->d[id].nv);
}
void orc_scene_get_verts(const OrcScene* s, int id, float* out) {
  memcpy(out, s->d[id].verts, sizeof(float) * 3 * s->d[id].nv);
}

/* ---- per-domain traversal ---- */
typedef struct { float t, u, v; uint32_t prim; int hit; } Hit;

static inline void tri_verts(const Domain* d, uint32_t prim, v3* v0, v3* v1, v3* v2) {
  const float* a = &d->verts[(size_t)d->faces[3 * (size_t)prim] * 3];
  const float* b = &d->verts[(size_t)d->faces[3 * (size_t)prim + 1] * 3];
  const float* c = &d->verts[(size_t)d->faces[3 * (size_t)prim + 2] * 3];
  *v0 = V(a[0], a[1], a[2]); *v1 = V(b[0], b[1], b[2]); *v2 = V(c[0], c[1], c[2]);
}

/* conservative slab test for the oracle's own BVH (not a reference function).  A zero direction
 * component is handled exactly: the slab passes iff lo <= o <= hi (inclusive), so a ray lying IN a box
 * face still visits the box. */
static inline void slab_axis(float lo, float hi, float o, float d, float inv, float* t0, float* t1) {
  if (d == 0.0f) {
    if (o < lo || o > hi) { *t0 = INFINITY; *t1 = -INFINITY; } else { *t0 = -INFINITY; *t1 = INFINITY; }
    return;
  }
  float a = (lo - o) * inv, b = (hi - o) * inv;
  *t0 = fminf(a, b); *t1 = fmaxf(a, b);
}
static inline int node_test(const BNode* n, v3 O, v3 D, v3 inv, float tbest, float* tn) {
  float x0, x1, y0, y1, z0, z1;
  slab_axis(n->lo[0], n->hi[0], O.x, D.x, inv.x, &x0, &x1);
  slab_axis(n->lo[1], n->hi[1], O.y, D.y, inv.y, &y0, &y1);
  slab_axis(n->lo[2], n->hi[2], O.z, D.z, inv.z, &z0, &z1);
  float tmin = fmaxf(fmaxf(x0, y0), z0);
  float tmax = fminf(fminf(x1, y1), z1);
  tmax = tmax + fabsf(tmax) * 1e-5f; /* slack: node boxes are exact FP32 vertex bounds, triangle t carries rounding */
  tmin = tmin - fabsf(tmin) * 1e-5f;
  *tn = tmin;
  return (tmin <= tmax) && (tmax >= 0.0f) && (tmin <= tbest);
}

/* closest: lexicographic min of (t, prim) over all passing triangles */
/* box tests only: a zero direction component would give 0*inf = NaN on a box face through the origin */
static inline v3 safe_inv(v3 D) {
  const float eps = 8.271806125530277e-25f; /* 2^-80 */
  float dx = fabsf(D.x) > eps ? D.x : copysignf(eps, D.x);
  float dy = fabsf(D.y) > eps ? D.y : copysignf(eps, D.y);
  float dz = fabsf(D.z) > eps ? D.z : copysignf(eps, D.z);
  return V(1.0f / dx, 1.0f / dy, 1.0f / dz);
}

static void domain_closest(const Domain* d, v3 O, v3 D, float tnear, Hit* h) {
  if (d->nnodes == 0) return;
  v3 inv = safe_inv(D);
  int stack[128]; int sp = 0; stack[sp++] = 0;
  while (sp) {
    const BNode* n = &d->nodes[stack[--sp]];
    float tn;
    if (!node_test(n, O, D, inv, h->hit ? h->t : INFINITY, &tn)) continue;
    if (n->count > 0) {
      for (int i = n->left; i < n->left + n->count; ++i) {
        uint32_t prim = d->prim[i];
        v3 v0, v1, v2; tri_verts(d, prim, &v0, &v1, &v2);
        float t, u, v;
        if (tri_test(O, D, v0, v1, v2, tnear, &t, &u, &v)) {
          if (!h->hit || t < h->t || (t == h->t && prim < h->prim)) {
            h->hit = 1; h->t = t; h->u = u; h->v = v; h->prim = prim;
          }
        }
      }
    } else {
      int l = n->left, r = -n->count;
      float tl, tr;
      int hl = node_test(&d->nodes[l], O, D, inv, h->hit ? h->t : INFINITY, &tl);
      int hr = node_test(&d->nodes[r], O, D, inv, h->hit ? h->t : INFINITY, &tr);
      if (hl && hr) { if (tl <= tr) { stack[sp++] = r; stack[sp++] = l; } else { stack[sp++] = l; stack[sp++] = r; } }
      else if (hl) stack[sp++] = l;
      else if (hr) stack[sp++] = r;
    }
  }
}

static int domain_any(const Domain* d, v3 O, v3 D, float tnear) {
  if (d->nnodes == 0) return 0;
  v3 inv = safe_inv(D);
  int stack[128]; int sp = 0; stack[sp++] = 0;
  while (sp) {
    const BNode* n = &d->nodes[stack[--sp]];
    float tn;
    if (!node_test(n, O, D, inv, INFINITY, &tn)) continue;
    if (n->count > 0) {
      for (int i = n->left; i < n->left + n->count; ++i) {
        v3 v0, v1, v2; tri_verts(d, d->prim[i], &v0, &v1, &v2);
        float t, u, v;
        if (tri_test(O, D, v0, v1, v2, tnear, &t, &u, &v)) return 1;
      }
    } else { stack[sp++] = -n->count; stack[sp++] = n->left; }
  }
  return 0;
}

/* ---- ray -> domain list: wbvh_embree.cc:126-140 + rays.h:71-79 ---- */
typedef struct { int id; float t; } DomHit;
static int cmp_domhit(const void* a, const void* b) {
  const DomHit* x = (const DomHit*)a; const DomHit* y = (const DomHit*)b;
  if (x->t < y->t) return -1;
  if (y->t < x->t) return 1;
  return x->id - y->id;
}
static int domain_list(const OrcScene* s, v3 O, v3 D, DomHit* out) {
  float org[3] = {O.x, O.y, O.z}, dir[3] = {D.x, D.y, D.z};
  int n = 0;
  for (int i = 0; i < s->nd; ++i) {
    float tmin, tmax;
    if (orc_intersect_aabb(s->d[i].bounds, org, dir, RAY_EPS, INFINITY, &tmin, &tmax)) {
      out[n].id = i; out[n].t = tmin; ++n;
    }
  }
  if (n > 1) qsort(out, n, sizeof(DomHit), cmp_domhit);
  return n;
}

static inline uint64_t make_key(float t, int dom, uint32_t prim) {
  return ((uint64_t)f2u(t) << 32) | ((uint64_t)(uint32_t)dom << 23) | (uint64_t)prim;
}

/* closest hit over all entered domains = what ooc::VBuf / insitu::VBuf converge to
 * (ooc_vbuf.cc:54-112, insitu_vbuf.h:144-152): min t; exact ties -> lowest (domain, prim)
 * (the reference is schedule-dependent there; documented near-tie set). */
typedef struct { int hit; float t, u, v; int dom; uint32_t prim; int ntests; } SceneHit;
static void scene_closest(const OrcScene* s, v3 O, v3 D, DomHit* scratch, SceneHit* out) {
  int n = domain_list(s, O, D, scratch);
  out->hit = 0; out->ntests = 0; out->t = INFINITY;
  for (int k = 0; k < n; ++k) {
    /* front-to-back early out is an optimisation only: a domain whose entry distance is beyond the
     * best hit (with slack) cannot hold a closer one */
    if (out->hit && scratch[k].t > out->t * 1.0001f + 1e-4f) break;
    Hit h; h.hit = 0; h.t = INFINITY; h.prim = 0; h.u = h.v = 0;
    domain_closest(&s->d[scratch[k].id], O, D, RAY_EPS, &h);
    out->ntests++;
    if (h.hit) {
      int id = scratch[k].id;
      if (!out->hit || h.t < out->t || (h.t == out->t && (id < out->dom || (id == out->dom && h.prim < out->prim)))) {
        out->hit = 1; out->t = h.t; out->u = h.u; out->v = h.v; out->dom = id; out->prim = h.prim;
      }
    }
  }
}
static int scene_occluded(const OrcScene* s, v3 O, v3 D, DomHit* scratch, int* ntests) {
  int n = domain_list(s, O, D, scratch);
  for (int k = 0; k < n; ++k) {
    if (ntests) (*ntests)++;
    if (domain_any(&s->d[scratch[k].id], O, D, RAY_EPS)) return 1;
  }
  return 0;
}

void orc_domain_lists(const OrcScene* s, const OrcRay* rays, size_t n, int cap, int* counts, int* ids,
                      float* tmins) {
#pragma omp parallel
  {
    DomHit* sc = (DomHit*)malloc(sizeof(DomHit) * (s->nd + 1));
#pragma omp for schedule(static)
    for (long long i = 0; i < (long long)n; ++i) {
      int c = domain_list(s, V(rays[i].ox, rays[i].oy, rays[i].oz), V(rays[i].dx, rays[i].dy, rays[i].dz), sc);
      counts[i] = c;
      for (int k = 0; k < c && k < cap; ++k) { ids[i * cap + k] = sc[k].id; tmins[i * cap + k] = sc[k].t; }
    }
    free(sc);
  }
}

void orc_intersect(const OrcScene* s, const OrcRay* rays, size_t n, uint64_t* keys, float* uv) {
#pragma omp parallel
  {
    DomHit* sc = (DomHit*)malloc(sizeof(DomHit) * (s->nd + 1));
#pragma omp for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i) {
      SceneHit h;
      scene_closest(s, V(rays[i].ox, rays[i].oy, rays[i].oz), V(rays[i].dx, rays[i].dy, rays[i].dz), sc, &h);
      keys[i] = h.hit ? make_key(h.t, h.dom, h.prim) : ORC_KEY_MISS;
      if (uv) { uv[2 * i] = h.hit ? h.u : 0.f; uv[2 * i + 1] = h.hit ? h.v : 0.f; }
    }
    free(sc);
  }
}

void orc_intersect_brute(const OrcScene* s, const OrcRay* rays, size_t n, uint64_t* keys, float* uv) {
#pragma omp parallel
  {
    DomHit* sc = (DomHit*)malloc(sizeof(DomHit) * (s->nd + 1));
#pragma omp for schedule(dynamic, 16)
    for (long long i = 0; i < (long long)n; ++i) {
      v3 O = V(rays[i].ox, rays[i].oy, rays[i].oz), D = V(rays[i].dx, rays[i].dy, rays[i].dz);
      int nd = domain_list(s, O, D, sc);
      uint64_t best = ORC_KEY_MISS; float bu = 0, bv = 0;
      for (int k = 0; k < nd; ++k) {
        const Domain* d = &s->d[sc[k].id];
        for (size_t p = 0; p < d->nf; ++p) {
          v3 v0, v1, v2; tri_verts(d, (uint32_t)p, &v0, &v1, &v2);
          float t, u, v;
          if (tri_test(O, D, v0, v1, v2, RAY_EPS, &t, &u, &v)) {
            uint64_t key = make_key(t, sc[k].id, (uint32_t)p);
            if (key < best) { best = key; bu = u; bv = v; }
          }
        }
      }
      keys[i] = best;
      if (uv) { uv[2 * i] = bu; uv[2 * i + 1] = bv; }
    }
    free(sc);
  }
}

/* single-domain query = Scene::intersect on one rtc_scene (render/scene.inl:190-199); the ray's own
 * tmin is honoured, tmax is ignored exactly like the reference (tfar = inf, rays.h:345-363) */
void orc_intersect_domain(const OrcScene* s, int dom, const OrcRay* rays, size_t n, uint64_t* keys, float* uv) {
#pragma omp parallel for schedule(dynamic, 256)
  for (long long i = 0; i < (long long)n; ++i) {
    Hit h; h.hit = 0; h.t = INFINITY; h.prim = 0; h.u = h.v = 0;
    domain_closest(&s->d[dom], V(rays[i].ox, rays[i].oy, rays[i].oz), V(rays[i].dx, rays[i].dy, rays[i].dz),
                   rays[i].tmin, &h);
    keys[i] = h.hit ? make_key(h.t, dom, h.prim) : ORC_KEY_MISS;
    if (uv) { uv[2 * i] = h.hit ? h.u : 0.f; uv[2 * i + 1] = h.hit ? h.v : 0.f; }
  }
}

void orc_occluded(const OrcScene* s, const OrcRay* rays, size_t n, uint8_t* occ) {
#pragma omp parallel
  {
    DomHit* sc = (DomHit*)malloc(sizeof(DomHit) * (s->nd + 1));
#pragma omp for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i)
      occ[i] = (uint8_t)scene_occluded(s, V(rays[i].ox, rays[i].oy, rays[i].oz),
                                       V(rays[i].dx, rays[i].dy, rays[i].dz), sc, NULL);
    free(sc);
  }
}
void orc_occluded_domain(const OrcScene* s, int dom, const OrcRay* rays, size_t n, uint8_t* occ) {
#pragma omp parallel for schedule(dynamic, 256)
  for (long long i = 0; i < (long long)n; ++i)
    occ[i] = (uint8_t)domain_any(&s->d[dom], V(rays[i].ox, rays[i].oy, rays[i].oz),
                                 V(rays[i].dx, rays[i].dy, rays[i].dz), rays[i].tmin);
}

/* ---- O1: double-precision truth ---- */
typedef struct { double t, u, v; int dom; int prim; } Cand;
static inline void consider(Cand* b1, Cand* b2, double t, double u, double v, int dom, int prim) {
  if (b1->dom < 0 || t < b1->t || (t == b1->t && (dom < b1->dom || (dom == b1->dom && prim < b1->prim)))) {
    *b2 = *b1; b1->t = t; b1->u = u; b1->v = v; b1->dom = dom; b1->prim = prim;
  } else if (b2->dom < 0 || t < b2->t) {
    b2->t = t; b2->u = u; b2->v = v; b2->dom = dom; b2->prim = prim;
  }
}
static void f64_domain(const Domain* d, int dom, v3 O, v3 D, int use_bvh, Cand* b1, Cand* b2) {
  d3 Od = D3(O), Dd = D3(D);
  double tnear = (double)RAY_EPS;
  if (!use_bvh) {
    for (size_t p = 0; p < d->nf; ++p) {
      v3 v0, v1, v2; tri_verts(d, (uint32_t)p, &v0, &v1, &v2);
      double t, u, v;
      if (tri_test_f64(Od, Dd, D3(v0), D3(v1), D3(v2), tnear, &t, &u, &v)) consider(b1, b2, t, u, v, dom, (int)p);
    }
    return;
  }
  if (d->nnodes == 0) return;
  const double deps = 1e-300;
  double ix = 1.0 / (fabs(Dd.x) > deps ? Dd.x : copysign(deps, Dd.x));
  double iy = 1.0 / (fabs(Dd.y) > deps ? Dd.y : copysign(deps, Dd.y));
  double iz = 1.0 / (fabs(Dd.z) > deps ? Dd.z : copysign(deps, Dd.z));
  int stack[128]; int sp = 0; stack[sp++] = 0;
  while (sp) {
    const BNode* n = &d->nodes[stack[--sp]];
    double lo3[3] = {n->lo[0], n->lo[1], n->lo[2]}, hi3[3] = {n->hi[0], n->hi[1], n->hi[2]};
    double o3[3] = {Od.x, Od.y, Od.z}, d3v[3] = {Dd.x, Dd.y, Dd.z}, i3[3] = {ix, iy, iz};
    double tmin = -INFINITY, tmax = INFINITY;
    for (int a = 0; a < 3; ++a) {
      double pad = 1e-6 * (fabs(lo3[a]) + fabs(hi3[a]) + 1e-3); /* admit FP32-graze candidates */
      if (d3v[a] == 0.0) {
        if (o3[a] < lo3[a] - pad || o3[a] > hi3[a] + pad) { tmin = INFINITY; tmax = -INFINITY; }
        continue;
      }
      double ta = (lo3[a] - pad - o3[a]) * i3[a], tb = (hi3[a] + pad - o3[a]) * i3[a];
      tmin = fmax(tmin, fmin(ta, tb)); tmax = fmin(tmax, fmax(ta, tb));
    }
    tmax = tmax * (1.0 + 1e-6) + 1e-9; tmin = tmin - fabs(tmin) * 1e-6 - 1e-9;
    double lim = b1->dom >= 0 ? b1->t * (1.0 + 1e-3) : INFINITY; /* keep everything within the tie window */
    if (!(tmin <= tmax) || !(tmax >= 0.0) || !(tmin <= lim)) continue;
    if (n->count > 0) {
      for (int i = n->left; i < n->left + n->count; ++i) {
        uint32_t p = d->prim[i];
        v3 v0, v1, v2; tri_verts(d, p, &v0, &v1, &v2);
        double t, u, v;
        if (tri_test_f64(Od, Dd, D3(v0), D3(v1), D3(v2), tnear, &t, &u, &v)) consider(b1, b2, t, u, v, dom, (int)p);
      }
    } else { stack[sp++] = -n->count; stack[sp++] = n->left; }
  }
}
void orc_intersect_f64(const OrcScene* s, const OrcRay* rays, size_t n, int use_bvh, int32_t* dom, int32_t* prim,
                       double* t, uint8_t* flags) {
#pragma omp parallel
  {
    DomHit* sc = (DomHit*)malloc(sizeof(DomHit) * (s->nd + 1));
#pragma omp for schedule(dynamic, 64)
    for (long long i = 0; i < (long long)n; ++i) {
      v3 O = V(rays[i].ox, rays[i].oy, rays[i].oz), D = V(rays[i].dx, rays[i].dy, rays[i].dz);
      int nd = domain_list(s, O, D, sc);
      Cand b1, b2; b1.dom = b2.dom = -1; b1.t = b2.t = INFINITY; b1.prim = b2.prim = -1; b1.u = b1.v = b2.u = b2.v = 0;
      for (int k = 0; k < nd; ++k) f64_domain(&s->d[sc[k].id], sc[k].id, O, D, use_bvh, &b1, &b2);
      dom[i] = b1.dom; prim[i] = b1.prim; t[i] = b1.dom >= 0 ? b1.t : INFINITY;
      uint8_t f = 0;
      if (b1.dom >= 0) {
        if (b2.dom >= 0 && fabs(b2.t - b1.t) <= 1e-5 * b1.t) f |= ORC_FLAG_NEAR_TIE;
        double w = 1.0 - b1.u - b1.v;
        double m = fmin(fmin(b1.u, b1.v), w);
        if (m < 1e-6) f |= ORC_FLAG_EDGE;
      }
      flags[i] = f;
    }
    free(sc);
  }
}

/* ---- hit attributes: trimesh_buffer.cc:328-360 ---- */
static void hit_attributes(const Domain* d, uint32_t prim, float u, float v, uint32_t* color, v3* Ns) {
  uint32_t i0 = d->faces[3 * (size_t)prim], i1 = d->faces[3 * (size_t)prim + 1], i2 = d->faces[3 * (size_t)prim + 2];
  *color = orc_interp_color(d->rgb[i0], d->rgb[i1], d->rgb[i2], u, v);
  float w = 1.f - u - v;
  const float* n0 = &d->normals[(size_t)i0 * 3]; const float* n1 = &d->normals[(size_t)i1 * 3];
  const float* n2 = &d->normals[(size_t)i2 * 3];
  Ns->x = (n0[0] * w) + (n1[0] * u) + (n2[0] * v);
  Ns->y = (n0[1] * w) + (n1[1] * u) + (n2[1] * v);
  Ns->z = (n0[2] * w) + (n1[2] * u) + (n2[2] * v);
}
void orc_hit_attributes(const OrcScene* s, const uint64_t* keys, const float* uv, size_t n, uint32_t* color,
                        float* ns3) {
  for (size_t i = 0; i < n; ++i) {
    if (keys[i] == ORC_KEY_MISS) { color[i] = 0; ns3[3 * i] = ns3[3 * i + 1] = ns3[3 * i + 2] = 0; continue; }
    int dom = (int)((keys[i] >> 23) & 0x1ff); uint32_t prim = (uint32_t)(keys[i] & 0x7fffff);
    v3 Ns; hit_attributes(&s->d[dom], prim, uv[2 * i], uv[2 * i + 1], &color[i], &Ns);
    ns3[3 * i] = Ns.x; ns3[3 * i + 1] = Ns.y; ns3[3 * i + 2] = Ns.z;
  }
}

/* ---- eye rays: ooc_tracer.inl:84-172 ---- */
/* eye_seed 1 = insitu::genMultiSampleEyeRays (insitu_ray.h:143-182): the jitter sampler is seeded with (pixid, s) */
void orc_gen_eye_rays_seed(const OrcCamera* cam, OrcTile tile, int spp, int eye_seed, OrcRay* rays);
void orc_gen_eye_rays(const OrcCamera* cam, OrcTile tile, int spp, OrcRay* rays) { orc_gen_eye_rays_seed(cam, tile, spp, 0, rays); }
void orc_gen_eye_rays_seed(const OrcCamera* cam, OrcTile tile, int spp, int eye_seed, OrcRay* rays) {
#pragma omp parallel for schedule(static)
  for (int y0 = 0; y0 < tile.h; ++y0) {
    for (int x0 = 0; x0 < tile.w; ++x0) {
      int x = tile.x + x0, y = tile.y + y0;
      for (int sidx = 0; sidx < spp; ++sidx) {
        int bufid = spp * (y0 * tile.w + x0) + sidx;
        float fx = (float)x, fy = (float)y;
        if (spp > 1) {
          uint32_t st = eye_seed ? orc_rng_init2(y * cam->image_w + x, sidx) : orc_rng_init1(bufid);
          fx = (float)x + orc_rng_get_float(&st);
          fy = (float)y + orc_rng_get_float(&st);
        }
        v3 d = cam_dir(cam, fx, fy);
        OrcRay* r = &rays[bufid];
        r->ox = cam->pos[0]; r->oy = cam->pos[1]; r->oz = cam->pos[2]; r->tmin = RAY_EPS;
        r->dx = d.x; r->dy = d.y; r->dz = d.z; r->tmax = INFINITY;
      }
    }
  }
}

/* ------------------------------------------------------------------------------------------------
 * Shading + ray-stream per sample.  ShaderPt: ooc/ooc_shader_pt.h:94-227 ; ShaderAo:
 * ooc/ooc_shader_ao.h:92-197 ; BSDFs: render/reflection.h:126-364 ; lights: render/light.h:41-89 ;
 * retire: ooc/ooc_tcontext.inl:123-135 + display/image.h:99-109.
 *
 * Speculation changes WHEN work is done, not the image: VBuf::update/correct
 * (ooc/ooc_vbuf.cc:41-112) keep exactly the closest hit's shadow/secondary rays.  The oracle therefore
 * walks each sample's ray chain non-speculatively.
 * ---------------------------------------------------------------------------------------------- */
static inline int has_positive(v3 v) { return v.x > 0.0f || v.y > 0.0f || v.z > 0.0f; }

static inline void image_add(float* fb, int pixid, v3 w, double scale) {
  /* buf.r += (scale * rgb): double product, float += double -> narrowed */
  fb[4 * (size_t)pixid + 0] = (float)((double)fb[4 * (size_t)pixid + 0] + scale * (double)w.x);
  fb[4 * (size_t)pixid + 1] = (float)((double)fb[4 * (size_t)pixid + 1] + scale * (double)w.y);
  fb[4 * (size_t)pixid + 2] = (float)((double)fb[4 * (size_t)pixid + 2] + scale * (double)w.z);
}

static float fr_dielectric(float cosThetaI, float etaI, float etaT, v3 wo, v3 n, v3* wt) {
  float sin2ThetaI = fmaxf(0.0f, 1.0f - (cosThetaI * cosThetaI));
  float eta = etaI / etaT;
  float sin2ThetaT = eta * eta * sin2ThetaI;
  if (sin2ThetaT >= 1.0f) return 1.0f;
  float cosThetaT = sqrtf(1.0f - sin2ThetaT);
  float Rparl = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
  float Rperp = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
  float fr = (Rparl * Rparl + Rperp * Rperp) / 2.0f;
  *wt = vadd(vscale(vneg(wo), eta), vscale(n, eta * cosThetaI - cosThetaT));
  return fr;
}
static int refract(float cosThetaI, float etaI, float etaT, v3 wo, v3 n, v3* wt) {
  float sin2ThetaI = fmaxf(0.0f, 1.0f - (cosThetaI * cosThetaI));
  float eta = etaI / etaT;
  float sin2ThetaT = eta * eta * sin2ThetaI;
  if (sin2ThetaT >= 1.0f) return 0;
  float cosThetaT = sqrtf(1.0f - sin2ThetaT);
  *wt = vadd(vscale(vneg(wo), eta), vscale(n, eta * cosThetaI - cosThetaT));
  return 1;
}

typedef struct { v3 org, dir, w; } Chain;

static void trace_sample(const OrcScene* s, const OrcShade* sh, v3 org, v3 dir, int pixid, int samid, float* fb,
                         DomHit* sc, OrcStats* st) {
  Chain stack[8]; int nstack = 0; int depth_of[8];
  stack[0].org = org; stack[0].dir = dir; stack[0].w = V(1.f, 1.f, 1.f); depth_of[0] = 0; nstack = 1;
  double scale = 1.0 / (double)sh->pixel_samples;
  v3 ks = V(sh->ks[0], sh->ks[1], sh->ks[2]);
  while (nstack) {
    Chain c = stack[--nstack]; int depth = depth_of[nstack];
    SceneHit h; scene_closest(s, c.org, c.dir, sc, &h);
    st->radiance_rays++; st->domain_tests_radiance += (uint64_t)h.ntests;
    if (!h.hit) continue;
    const Domain* d = &s->d[h.dom];
    uint32_t color; v3 normal; hit_attributes(d, h.prim, h.u, h.v, &color, &normal);
    v3 pos = V(c.dir.x * h.t + c.org.x, c.dir.y * h.t + c.org.y, c.dir.z * h.t + c.org.z);
    float kdv[3]; orc_unpack_color(color, kdv); v3 kd = V(kdv[0], kdv[1], kdv[2]);
    v3 wo = vneg(c.dir);
    v3 Lin = c.w;
    float cos_theta_i = gdot(wo, normal);
    int entering = cos_theta_i > 0.0f;
    v3 nff = gnormalize(entering ? normal : vneg(normal));
    int delta = d->mtl_type != ORC_MTL_DIFFUSE;
    int next_depth = depth + 1; /* next_actual_depth = ray_depth + rayin.depth + 1 */
    v3 wi, Lr; float pdf, costheta;
    if (sh->shader == 0) {
      if (!delta) {
        uint32_t ls = orc_rng_init1((int)((uint32_t)samid * (uint32_t)next_depth));
        for (int l = 0; l < sh->nlights; ++l) {
          const OrcLight* L = &sh->lights[l];
          v3 rad = V(L->radiance[0], L->radiance[1], L->radiance[2]);
          if (L->type == 1) {
            for (int k = 0; k < sh->ao_samples; ++k) {
              float u1 = orc_rng_get_float(&ls), u2 = orc_rng_get_float(&ls);
              cosine_hemisphere(u1, u2, nff, &wi, &pdf);
              if (pdf > 0.0f) {
                costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
                Lr = vscale(vmul(Lin, blinn_phong(costheta, kd, ks, sh->shininess, rad, wi, nff, wo)),
                            1.0f / (pdf * (float)sh->ao_samples));
                if (has_positive(Lr)) {
                  int nt = 0; st->shadow_rays++;
                  if (!scene_occluded(s, pos, wi, sc, &nt)) image_add(fb, pixid, Lr, scale);
                  st->domain_tests_shadow += (uint64_t)nt;
                }
              }
            }
          } else {
            wi = gnormalize(vsub(V(L->position[0], L->position[1], L->position[2]), pos));
            pdf = 1.0f;
            costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
            Lr = vscale(vmul(Lin, blinn_phong(costheta, kd, ks, sh->shininess, rad, wi, nff, wo)), 1.0f / pdf);
            if (has_positive(Lr)) {
              int nt = 0; st->shadow_rays++;
              if (!scene_occluded(s, pos, wi, sc, &nt)) image_add(fb, pixid, Lr, scale);
              st->domain_tests_shadow += (uint64_t)nt;
            }
          }
        }
      }
    } else {
      float ao_weight = 1.0f / (float)sh->ao_samples;
      for (int l = 0; l < sh->ao_samples; ++l) {
        uint32_t ls = orc_rng_init1((int)((uint32_t)pixid * (uint32_t)(l + 1)));
        float u1 = orc_rng_get_float(&ls), u2 = orc_rng_get_float(&ls);
        if (delta) break; /* reference: LOG(FATAL) "forbidden" (reflection.h:253-256) */
        cosine_hemisphere(u1, u2, nff, &wi, &pdf);
        costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
        Lr = vscale(vmul(Lin, kd), ONE_OVER_PI * costheta * ao_weight / pdf);
        if (has_positive(Lr)) {
          int nt = 0; st->shadow_rays++;
          if (!scene_occluded(s, pos, wi, sc, &nt)) image_add(fb, pixid, Lr, scale);
          st->domain_tests_shadow += (uint64_t)nt;
        }
      }
    }
    if (next_depth < sh->bounces) {
      wo = gnormalize(wo);
      if (delta) {
        if (cos_theta_i != 0.0f) {
          cos_theta_i = gclamp(cos_theta_i, -1.0f, 1.0f);
          float abs_cos = fabsf(cos_theta_i);
          if (!entering) cos_theta_i = abs_cos;
          int has_reflect = 0, has_trans = 0; float fr = 0.f; v3 wt = V(0, 0, 0);
          if (d->mtl_type == ORC_MTL_MIRROR) { fr = 1.0f; has_reflect = 1; }
          else if (d->mtl_type == ORC_MTL_GLASS) {
            float etaI = entering ? d->mtl[0] : d->mtl[1], etaT = entering ? d->mtl[1] : d->mtl[0];
            fr = fr_dielectric(cos_theta_i, etaI, etaT, wo, nff, &wt);
            if (fr == 1.0f) has_reflect = 1; else if (fr == 0.0f) has_trans = 1; else { has_reflect = 1; has_trans = 1; }
          } else {
            float etaI = entering ? d->mtl[0] : d->mtl[1], etaT = entering ? d->mtl[1] : d->mtl[0];
            if (refract(cos_theta_i, etaI, etaT, wo, nff, &wt)) { has_trans = 1; fr = 0.0f; }
            else { has_reflect = 1; fr = 1.0f; }
          }
          /* reference CHECK-fails when glass yields both (ooc_shader_pt.h:199-201); we emit both, in the
           * reference's push order (reflection first) */
          if (has_trans && nstack < 8) {
            wi = gnormalize(wt);
            Lr = vscale(Lin, (1.0f - fr) / abs_cos);
            if (has_positive(Lr)) { stack[nstack].org = pos; stack[nstack].dir = wi; stack[nstack].w = Lr; depth_of[nstack] = next_depth; nstack++; }
          }
          if (has_reflect && nstack < 8) {
            v3 wr = vadd(vneg(wo), vscale(nff, 2.0f * gdot(wo, nff)));
            wi = gnormalize(wr);
            Lr = vscale(Lin, fr / abs_cos);
            if (has_positive(Lr)) { stack[nstack].org = pos; stack[nstack].dir = wi; stack[nstack].w = Lr; depth_of[nstack] = next_depth; nstack++; }
          }
        }
      } else {
        uint32_t rs = orc_rng_init1((int)((uint32_t)samid * (uint32_t)next_depth));
        float u1 = orc_rng_get_float(&rs), u2 = orc_rng_get_float(&rs);
        cosine_hemisphere(u1, u2, nff, &wi, &pdf);
        costheta = gclamp(gdot(nff, wi), 0.0f, 1.0f);
        /* Lin * surf_radiance * ONE_OVER_PI * costheta / pdf : left to right */
        Lr = vscale(vscale(vmul(Lin, kd), ONE_OVER_PI), costheta);
        Lr = V(Lr.x / pdf, Lr.y / pdf, Lr.z / pdf);
        if (has_positive(Lr) && nstack < 8) {
          stack[nstack].org = pos; stack[nstack].dir = wi; stack[nstack].w = Lr; depth_of[nstack] = next_depth; nstack++;
        }
      }
    }
  }
}

void orc_render_tile(const OrcScene* s, const OrcCamera* cam, const OrcShade* sh, OrcTile tile, float* fb,
                     OrcStats* stats) {
  OrcStats total; memset(&total, 0, sizeof(total));
  int spp = sh->pixel_samples;
  long long nsamples = (long long)tile.w * tile.h * spp;
#pragma omp parallel
  {
    DomHit* sc = (DomHit*)malloc(sizeof(DomHit) * (s->nd + 1));
    OrcStats st; memset(&st, 0, sizeof(st));
    /* one pixel per iteration so that all samples of a pixel are added by one thread (the reference
     * serialises image adds with an omp critical, ooc_pcontext.h:165-166) */
#pragma omp for schedule(dynamic, 64)
    for (long long p = 0; p < (long long)tile.w * tile.h; ++p) {
      int x0 = (int)(p % tile.w), y0 = (int)(p / tile.w);
      int x = tile.x + x0, y = tile.y + y0;
      int pixid = y * cam->image_w + x;
      for (int sidx = 0; sidx < spp; ++sidx) {
        int bufid = spp * (y0 * tile.w + x0) + sidx;
        float fx = (float)x, fy = (float)y;
        if (spp > 1) {
          uint32_t rs = sh->eye_seed ? orc_rng_init2(pixid, sidx) : orc_rng_init1(bufid);
          fx = (float)x + orc_rng_get_float(&rs);
          fy = (float)y + orc_rng_get_float(&rs);
        }
        v3 dir = cam_dir(cam, fx, fy);
        st.eye_rays++;
        trace_sample(s, sh, V(cam->pos[0], cam->pos[1], cam->pos[2]), dir, pixid, bufid, fb, sc, &st);
      }
    }
#pragma omp critical
    {
      total.eye_rays += st.eye_rays; total.radiance_rays += st.radiance_rays; total.shadow_rays += st.shadow_rays;
      total.domain_tests_radiance += st.domain_tests_radiance; total.domain_tests_shadow += st.domain_tests_shadow;
    }
    free(sc);
  }
  (void)nsamples;
  if (stats) {
    stats->eye_rays += total.eye_rays; stats->radiance_rays += total.radiance_rays;
    stats->shadow_rays += total.shadow_rays; stats->domain_tests_radiance += total.domain_tests_radiance;
    stats->domain_tests_shadow += total.domain_tests_shadow;
  }
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* ---- counter-based triangle soup (bench configs C3 / C4) -------------------------------------------------
 * CPU restatement of the device generator (spray_b200/csrc/build.cu, soup_kernel): Philox-4x32-10 (Salmon, Moraes,
 * Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11; constants and round function as published, known
 * answers in tests/test_oracle_kat.py), key (key, 0), counter (primID, block, 0, 0).  IEEE-exact float operations only
 * (this file is built with -ffp-contract=off), so the two sides produce the same bits. */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t x0 = ctr[0], x1 = ctr[1], x2 = ctr[2], x3 = ctr[3], ka = key[0], kb = key[1];
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * x0, p1 = (uint64_t)0xCD9E8D57u * x2;
    const uint32_t y0 = (uint32_t)(p1 >> 32) ^ x1 ^ ka, y2 = (uint32_t)(p0 >> 32) ^ x3 ^ kb;
    x0 = y0; x1 = (uint32_t)p1; x2 = y2; x3 = (uint32_t)p0;
    ka += 0x9E3779B9u; kb += 0xBB67AE85u;
  }
  out[0] = x0; out[1] = x1; out[2] = x2; out[3] = x3;
}
static inline float soup_u01(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-08f; }
static void soup_edge(uint32_t key, uint32_t prim, uint32_t block0, float len, float e[3]) {
  float dx = 1.0f, dy = 0.0f, dz = 0.0f;
  const uint32_t k2[2] = {key, 0u};
  for (uint32_t k = 0; k < 16u; ++k) {
    const uint32_t c[4] = {prim, block0 + k, 0u, 0u};
    uint32_t r[4];
    orc_philox4x32_10(c, k2, r);
    const float x = 2.0f * soup_u01(r[0]) - 1.0f, y = 2.0f * soup_u01(r[1]) - 1.0f, z = 2.0f * soup_u01(r[2]) - 1.0f;
    const float q = x * x + y * y + z * z;
    if (q > 1.0e-4f && q <= 1.0f) {
      const float inv = 1.0f / sqrtf(q);
      dx = x * inv; dy = y * inv; dz = z * inv;
      break;
    }
  }
  e[0] = dx * len; e[1] = dy * len; e[2] = dz * len;
}
void orc_gen_soup(uint32_t key, size_t ntris, const float lo[3], const float hi[3], float h, float* xyz, uint32_t* idx,
                  uint32_t* rgb) {
  const uint32_t k2[2] = {key, 0u};
#pragma omp parallel for schedule(static)
  for (int64_t pp = 0; pp < (int64_t)ntris; ++pp) {
    const uint32_t p = (uint32_t)pp;
    const uint32_t c0[4] = {p, 0u, 0u, 0u}, c1[4] = {p, 1u, 0u, 0u};
    uint32_t r0[4], r1[4];
    orc_philox4x32_10(c0, k2, r0);
    orc_philox4x32_10(c1, k2, r1);
    float c[3], e1[3], e2[3];
    for (int a = 0; a < 3; ++a) c[a] = lo[a] + soup_u01(r0[a]) * (hi[a] - lo[a]);
    const float len1 = (0.5f + 1.5f * soup_u01(r0[3])) * h, len2 = (0.5f + 1.5f * soup_u01(r1[0])) * h;
    soup_edge(key, p, 2u, len1, e1);
    soup_edge(key, p, 18u, len2, e2);
    float* v = xyz + 9 * (size_t)p;
    for (int a = 0; a < 3; ++a) {
      v[a] = fminf(fmaxf(c[a], lo[a]), hi[a]);
      v[3 + a] = fminf(fmaxf(c[a] + e1[a], lo[a]), hi[a]);
      v[6 + a] = fminf(fmaxf(c[a] + e2[a], lo[a]), hi[a]);
    }
    for (uint32_t k = 0; k < 3u; ++k) {
      const uint32_t i = 3u * p + k;
      if (idx) idx[i] = i;
      if (rgb) {
        const uint32_t hsh = (i * 2654435761u) ^ (i >> 13);
        rgb[i] = ((64u + (hsh & 0x7Fu)) << 16) | ((64u + ((hsh >> 8) & 0x7Fu)) << 8) | (64u + ((hsh >> 16) & 0x7Fu));
      }
    }
  }
}
