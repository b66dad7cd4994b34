/*
 * spray_oracle.h — CPU ORACLE for the per-domain ray-tracing hot path of hyungman/SpRay.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product (spray_b200/, host/)
 * never includes, links or calls anything in this directory.
 *
 * PARITY STATUS: **parity unpinned by the reference** — the reference repository has no tests, no
 * golden vectors and no numeric fixtures for this path, and its ray/triangle
 * arithmetic lives in Embree 2.x (un-vendored, "tested with Embree v2.17.1", docs/build_linux.md:3),
 * which is absent from this container and cannot be built (no network).  The reference itself cannot be
 * compiled here either (needs Embree 2, MPI, IceT, OpenGL).  What *can* be pinned is pinned by
 * known-answer tests on the reference's own closed-form code (RNG streams, Morton codes, slab test,
 * colour truncation, tile lists, LRU sequences) in tests/test_oracle_kat.py, and by the INPUT fixtures
 * under tests/golden/ generated from the reference's example files.
 *
 * Every function cites the reference file:line it restates (paths relative to /root/reference).
 *
 * Three tracers:
 *   O1  orc_intersect_f64     double precision, all triangles of all AABB-entered domains -> truth
 *                             (domain, primID, t) + near-tie / edge-graze flags
 *   O2  orc_intersect/occluded/render_tile   FP32 tracer with the *specified* FP32 operation order
 *                             (see "FP32 contract" in DESIGN.md) over a binned-SAH BVH2; OpenMP.
 *                             This is the timed CPU baseline ("port", not Embree).
 *   O2b orc_intersect_brute   same FP32 triangle test, no BVH -> validates O2's culling.
 */
#ifndef SPRAY_ORACLE_H
#define SPRAY_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float ox, oy, oz, tmin, dx, dy, dz, tmax; } OrcRay; /* 32 B, = Ray32 of the C ABI */

typedef struct {
  float pos[3];
  float lowerleft[3];
  float wvec[3];
  float hvec[3];
  int image_w, image_h;
} OrcCamera;

typedef struct { int x, y, w, h; } OrcTile;

typedef struct {
  int type; /* 0 = point (light.h:41-63), 1 = diffuse hemisphere (light.h:65-89) */
  float position[3];
  float radiance[3];
} OrcLight;

typedef struct {
  int shader;        /* 0 = ShaderPt (ooc_shader_pt.h), 1 = ShaderAo (ooc_shader_ao.h) */
  int bounces;       /* cfg.bounces */
  int pixel_samples; /* cfg.pixel_samples */
  int ao_samples;    /* cfg.ao_samples */
  float ks[3];
  float shininess;
  int nlights;
  const OrcLight* lights;
  int eye_seed;      /* 0: ooc tracers, RandomSampler_init(bufid) (ooc_tracer.inl:151-155);
                        1: in-situ tracers, RandomSampler_init(pixid, s) (insitu_ray.h:161-165) */
} OrcShade;

typedef struct {
  uint64_t eye_rays, radiance_rays, shadow_rays, domain_tests_radiance, domain_tests_shadow;
} OrcStats;

enum { ORC_MTL_DIFFUSE = 0, ORC_MTL_MIRROR = 1, ORC_MTL_GLASS = 2, ORC_MTL_TRANSMISSION = 3 };
#define ORC_KEY_MISS 0xFFFFFFFFFFFFFFFFull
#define ORC_FLAG_NEAR_TIE 1
#define ORC_FLAG_EDGE 2

typedef struct OrcScene OrcScene;

/* ---- closed-form pieces (known-answer tests) ---- */
uint32_t orc_rng_init1(int id);
uint32_t orc_rng_init2(int pixel_id, int sample_id);
float orc_rng_get_float(uint32_t* state);
uint32_t orc_morton(float x, float y, float z);
int orc_intersect_aabb(const float bounds6[6], const float org[3], const float dir[3], float t0, float t1,
                       float* tmin_o, float* tmax_o);
uint32_t orc_interp_color(uint32_t c0, uint32_t c1, uint32_t c2, float u, float v);
void orc_unpack_color(uint32_t c, float rgb[3]);
void orc_camera_init(const float campos[3], const float lookat[3], const float up[3], float vfov, int image_w,
                     int image_h, OrcCamera* out);
void orc_default_camera_pose(const float scene_bounds6[6], float campos[3], float lookat[3], float up[3]);
void orc_camera_ray(const OrcCamera* cam, float x, float y, float org[3], float dir[3]);
int orc_image_schedule_tiles(int64_t image_w, int64_t image_h, int64_t spp, int64_t num_ranks, int rank,
                             int64_t max_samples_per_rank, OrcTile* out, int cap);
void orc_partition(int ndomains, const float* world_bounds6, const float scene_bounds6[6], int num_ranks,
                   int* domain_to_rank, uint32_t* codes);
void orc_cosine_hemisphere(float u1, float u2, const float n[3], float wi[3], float* pdf);
void orc_blinn_phong(float costheta, const float kd[3], const float ks[3], float shininess, const float li[3],
                     const float wi[3], const float n[3], const float wo[3], float out[3]);
void orc_ppm_quantize(const float* rgba, int n, uint32_t* out_rgb);
void orc_tri_test(const float org[3], const float dir[3], const float v0[3], const float v1[3], const float v2[3],
                  float tnear, int* hit, float* t, float* u, float* v);
/* LRU domain cache (lru_cache.cc:65-171): returns hit(1)/miss(0) sequence + block ids */
void orc_lru_sequence(int ndomains, int capacity, const int* requests, int n, int* hits, int* blocks);

/* ---- scene ---- */
OrcScene* orc_scene_create(int ndomains);
void orc_scene_destroy(OrcScene*);
/* verts are OBJECT space; if mat4 != NULL (column-major 16 floats) it is applied as
 * trimesh_buffer.cc:141-157 does; rgb packed 0x00RRGGBB per vertex (may be NULL -> 0). */
void orc_scene_set_domain(OrcScene*, int id, const float* verts, size_t nv, const uint32_t* faces, size_t nf,
                          const uint32_t* rgb, const float* mat4, const float world_bounds6[6], int mtl_type,
                          const float mtl_params[3]);
void orc_scene_commit(OrcScene*);
void orc_scene_bounds(const OrcScene*, float out6[6]);
size_t orc_scene_num_verts(const OrcScene*, int id);
void orc_scene_get_normals(const OrcScene*, int id, float* out /* nv*3 */);
void orc_scene_get_verts(const OrcScene*, int id, float* out /* nv*3, world space */);

/* ---- rays ---- */
void orc_gen_eye_rays(const OrcCamera* cam, OrcTile tile, int spp, OrcRay* rays /* tile.w*tile.h*spp */);
/* per ray: all entered domains sorted by (tmin,id) (wbvh_embree.cc:126-140, rays.h:71-79) */
void orc_domain_lists(const OrcScene*, const OrcRay* rays, size_t n, int cap, int* counts, int* ids,
                      float* tmins);
/* key = (float_bits(t)<<32) | (domain<<23) | prim ; ORC_KEY_MISS if no hit */
void orc_intersect(const OrcScene*, const OrcRay* rays, size_t n, uint64_t* keys, float* uv);
void orc_intersect_brute(const OrcScene*, const OrcRay* rays, size_t n, uint64_t* keys, float* uv);
void orc_intersect_domain(const OrcScene*, int domain, const OrcRay* rays, size_t n, uint64_t* keys, float* uv);
void orc_occluded(const OrcScene*, const OrcRay* rays, size_t n, uint8_t* occ);
void orc_occluded_domain(const OrcScene*, int domain, const OrcRay* rays, size_t n, uint8_t* occ);
void orc_intersect_f64(const OrcScene*, const OrcRay* rays, size_t n, int use_bvh, int32_t* dom, int32_t* prim,
                       double* t, uint8_t* flags);
/* hit attributes (trimesh_buffer.cc:328-360): packed colour + unnormalised Ns */
void orc_hit_attributes(const OrcScene*, const uint64_t* keys, const float* uv, size_t n, uint32_t* color,
                        float* ns3);

/* ---- full path: eye rays -> closest hit -> shade -> shadow/bounce -> framebuffer ---- */
void orc_gen_eye_rays_seed(const OrcCamera* cam, OrcTile tile, int spp, int eye_seed, OrcRay* rays);
void orc_render_tile(const OrcScene*, const OrcCamera*, const OrcShade*, OrcTile tile, float* fb_rgba,
                     OrcStats* stats);
int orc_num_threads(void);
/* counter-based triangle soup of bench configs C3 / C4 (restates the device generator bit for bit) */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void orc_gen_soup(uint32_t key, size_t ntris, const float lo[3], const float hi[3], float h, float* xyz, uint32_t* idx,
                  uint32_t* rgb);

#ifdef __cplusplus
}
#endif
#endif
