/*
 * spray_gpu.h — C ABI of libspraygpu: the B200 (sm_100a) replacement for the Embree calls on SpRay's
 * per-domain ray-tracing hot path.  Plain pointers and sizes only; no C++/torch types.
 *
 * Every entry point names the reference interface it replaces (paths relative to the SpRay tree).
 * All functions return SPRAYGPU_OK (0) or a negative error code; spraygpu_last_error() gives the text.
 * The reference never propagates errors (CHECK/LOG(FATAL) abort, trimesh_buffer.cc:68-113); the C++
 * shim in host/ aborts on non-zero to mirror that.
 *
 * Pointer arguments marked [h|d] may be host or device pointers (detected with
 * cudaPointerGetAttributes).  Host pointers are staged through device memory inside the call and the
 * call returns after the results are back on the host; device pointers are used in place, the work is
 * enqueued on the device's stream and the call returns without synchronising.
 *
 * Threading: intersect/occluded/trace calls on committed scenes may be issued from any host thread
 * but are serialised per spraygpu_device (one stream).  set_mesh/commit/destroy must be externally
 * serialised per scene — the same contract as the reference (load/commit inside `omp single`,
 * src/ooc/ooc_pcontext.h:145-154).
 */
#ifndef SPRAY_GPU_H
#define SPRAY_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SPRAYGPU_API __attribute__((visibility("default")))
#else
#define SPRAYGPU_API
#endif

#define SPRAYGPU_OK 0
#define SPRAYGPU_ERR_CUDA (-1)
#define SPRAYGPU_ERR_ARG (-2)
#define SPRAYGPU_ERR_STATE (-3)
#define SPRAYGPU_ERR_LIMIT (-4)

#define SPRAYGPU_MAX_DOMAINS 512      /* 9 key bits */
#define SPRAYGPU_MAX_PRIMS (1u << 23) /* 23 key bits */
#define SPRAYGPU_MAX_LIGHTS 8
#define SPRAYGPU_KEY_MISS 0xFFFFFFFFFFFFFFFFull

typedef struct spraygpu_device_s* spraygpu_device;
typedef struct spraygpu_scene_s* spraygpu_scene;

/* Ray record, 32 B.  Replaces the org/dir/tnear/tfar head of RTCRay / RTCRayIntersection
 * (src/render/rays.h:117-170,246-274).  tmin is honoured; tmax is IGNORED (= inf) exactly like every
 * reference call site (rays.h:345-363,389-423 set tfar = SPRAY_FLOAT_INF for radiance AND shadow).
 * DEVICE arrays of rays must be 16-byte aligned (any cudaMalloc / torch allocation is); host arrays need not be. */
typedef struct { float ox, oy, oz, tmin, dx, dy, dz, tmax; } spraygpu_ray;

/* Hit key: (float_bits(t) << 32) | (domain << 23) | primID ; SPRAYGPU_KEY_MISS when nothing was hit.
 * Monotone in t because t > tnear = 0.001 > 0 (src/render/spray.h:46).  Replaces the
 * tfar/geomID/primID output fields of RTCRay and the per-sample t-buffer entry of ooc::VBuf /
 * insitu::VBuf (src/ooc/ooc_vbuf.cc:54-112, src/insitu/insitu_vbuf.h:74-152). */
typedef uint64_t spraygpu_key;

enum { SPRAYGPU_MTL_DIFFUSE = 0, SPRAYGPU_MTL_MIRROR = 1, SPRAYGPU_MTL_GLASS = 2, SPRAYGPU_MTL_TRANSMISSION = 3 };
/* Domain material: src/render/domain.h:34 (Bsdf*), src/io/scene_loader.cc:89-141.
 * params: diffuse albedo rgb | mirror reflectance rgb | glass/transmission eta_exterior, eta_interior. */
typedef struct { int32_t type; float params[3]; } spraygpu_material;

enum { SPRAYGPU_LIGHT_POINT = 0, SPRAYGPU_LIGHT_DIFFUSE = 1 };
/* src/render/light.h:41-89 ; src/io/scene_loader.cc:199-228 */
typedef struct { int32_t type; float position[3]; float radiance[3]; } spraygpu_light;

/* Pinhole camera frame as src/render/camera.h:128-160 stores it (computed on the host). */
typedef struct {
  float pos[3], lowerleft[3], wvec[3], hvec[3];
  int32_t image_w, image_h;
} spraygpu_camera;

/* A blocking tile (src/render/tile.h Tile; produced by ImageScheduleTileList, tile.cc:317-391).
 * samid of sample s of pixel (x,y) = spp * ((y - tile.y) * tile.w + (x - tile.x)) + s
 * (src/ooc/ooc_tracer.inl:84-172). */
typedef struct { int32_t x, y, w, h; } spraygpu_tile;

enum { SPRAYGPU_SHADER_PT = 0, SPRAYGPU_SHADER_AO = 1 };
/* Shader configuration: ShaderPt::init / ShaderAo::init (src/ooc/ooc_shader_pt.h:45-52,
 * ooc_shader_ao.h) + Config fields (src/render/config.cc:33-73). */
typedef struct {
  int32_t shader;         /* SPRAYGPU_SHADER_* */
  int32_t bounces;        /* cfg.bounces (>= 1) */
  int32_t pixel_samples;  /* cfg.pixel_samples */
  int32_t ao_samples;     /* cfg.ao_samples (area-light / AO sample count) */
  float ks[3];            /* cfg.ks   (--blinn) */
  float shininess;        /* cfg.shininess */
  int32_t nlights;
  spraygpu_light lights[SPRAYGPU_MAX_LIGHTS];
  int32_t eye_seed;       /* SPRAYGPU_EYE_SEED_*: how multi-sample eye rays are jittered (ignored at 1 spp) */
} spraygpu_shade;
/* The reference's two tracer families seed the pixel jitter differently:
 *   OOC    RandomSampler_init(sampler, bufid)     bufid = spp * tile-relative pixel + s   (src/ooc/ooc_tracer.inl:151-155)
 *   INSITU RandomSampler_init(sampler, pixid, s)  pixid = image_w * y + x                 (src/insitu/insitu_ray.h:161-165)
 * spraygpu_trace_tile_insitu always uses INSITU; the other trace calls follow shade->eye_seed. */
enum { SPRAYGPU_EYE_SEED_OOC = 0, SPRAYGPU_EYE_SEED_INSITU = 1 };

typedef struct {
  uint64_t eye_rays;        /* primary rays generated */
  uint64_t radiance_rays;   /* closest-hit rays resolved (primary + bounce) */
  uint64_t shadow_rays;     /* any-hit rays resolved */
  uint64_t domain_tests;    /* (ray, domain) traversals started */
  uint64_t nodes_visited;   /* BVH8 nodes fetched  (counters build only, else 0) */
  uint64_t tris_tested;     /* triangles fetched   (counters build only, else 0) */
  uint64_t hits;            /* radiance rays that hit */
} spraygpu_stats;

typedef struct {
  uint64_t num_vertices, num_faces, num_nodes;
  uint64_t node_bytes, tri_bytes, attr_bytes;
  float bounds[6];
  double build_ms;  /* device time of the last commit (CUDA events) */
  uint32_t bvh_depth; /* levels of the 8-wide BVH; commit fails with SPRAYGPU_ERR_LIMIT above the traversal stack (48) */
  uint32_t reserved;
} spraygpu_scene_info;

SPRAYGPU_API const char* spraygpu_last_error(void);
SPRAYGPU_API const char* spraygpu_version(void);

/* ---- device: replaces rtcNewDevice / rtcDeleteDevice (src/render/trimesh_buffer.cc:104,178;
 *      wbvh_embree.cc:60) ---- */
SPRAYGPU_API int spraygpu_device_create(int cuda_device, spraygpu_device* out);
SPRAYGPU_API int spraygpu_device_destroy(spraygpu_device dev);
/* Launch on an existing cudaStream_t (e.g. torch's current stream) instead of the default stream. */
SPRAYGPU_API int spraygpu_device_set_stream(spraygpu_device dev, void* cuda_stream);
SPRAYGPU_API int spraygpu_device_synchronize(spraygpu_device dev);
/* kernels launched by this library since device_create (bench.py's gpu_launches) */
SPRAYGPU_API uint64_t spraygpu_device_launch_count(spraygpu_device dev);

/* ---- per-domain scene: replaces rtcDeviceNewScene / rtcDeleteScene / rtcNewTriangleMesh /
 *      rtcSetBuffer2 / rtcUpdate / rtcEnable / rtcCommit (src/render/trimesh_buffer.cc:112,174,189-225) ---- */
SPRAYGPU_API int spraygpu_scene_create(spraygpu_device dev, spraygpu_scene* out);
SPRAYGPU_API int spraygpu_scene_destroy(spraygpu_scene scene);
/* COPIES the mesh to the device (Embree borrows host arrays via rtcSetBuffer2; a GPU cannot).
 * xyz: nv*3 floats (object space), idx: nf*3 uint32, rgb: nv packed 0x00RRGGBB (src/io/ply_loader.cc:247)
 * or NULL.  [h|d] */
SPRAYGPU_API int spraygpu_scene_set_mesh(spraygpu_scene scene, const float* xyz, size_t nv, const uint32_t* idx, size_t nf,
                            const uint32_t* rgb);
/* Column-major 4x4 applied to every vertex at commit with glm's mat4*vec4 operation order
 * (TriMeshBuffer::load, src/render/trimesh_buffer.cc:141-157).  NULL = identity (no transform pass). */
SPRAYGPU_API int spraygpu_scene_set_transform(spraygpu_scene scene, const float* mat4_colmajor);
/* transform -> per-vertex normals (TriMeshBuffer::computeNormals, trimesh_buffer.cc:267-326, bit-exact:
 * same per-vertex summation order) -> LBVH -> 8-wide compressed BVH -> attribute pack.  = rtcCommit. */
SPRAYGPU_API int spraygpu_scene_commit(spraygpu_scene scene);
/* Parity dump of the acceleration structure (Embree's is opaque; here the tests compare it node for node with the CPU
 * emulation of the same builder): num_nodes x 80-byte compressed 8-wide nodes and num_faces x 48-byte triangle records
 * in leaf order.  Either pointer may be NULL. */
SPRAYGPU_API int spraygpu_scene_get_bvh(spraygpu_scene scene, void* nodes, void* tris);
/* The same build, ENQUEUED only: the whole pipeline (sort, hierarchy, collapse, records) runs on the device without a
 * host round trip, so the caller can upload the next domain's mesh while this one builds (the reference serialises
 * PLY load and rtcCommit of every cache miss behind a barrier, src/ooc/ooc_pcontext.h:145-154).  The scene may be
 * bound (spraygpu_domain_bind) right away; node count, depth check and bounds are read back the first time they are
 * needed (spraygpu_scene_get_info, or the next trace call over the domain table, which fails if the build did). */
SPRAYGPU_API int spraygpu_scene_commit_async(spraygpu_scene scene);
SPRAYGPU_API int spraygpu_scene_get_info(spraygpu_scene scene, spraygpu_scene_info* out);
SPRAYGPU_API int spraygpu_scene_get_normals(spraygpu_scene scene, float* out_nv3);      /* [h|d] */
SPRAYGPU_API int spraygpu_scene_get_vertices(spraygpu_scene scene, float* out_nv3);     /* world space, [h|d] */

/* The benchmark's synthetic domains ("triangle soup, generated on device from a counter-based
 * RNG ... reproducible on the CPU oracle from the same function"); stands in for PlyLoader::load of a mesh file
 * (src/io/ply_loader.cc:179-324) when there is no file.  Triangle p of the soup is a pure function of (key, p):
 * Philox-4x32-10 with key (key, 0) and counter (p, block, 0, 0); centre uniform in the cell [lo, hi], two edges of uniform
 * direction and length U(0.5, 2) * h, vertices clipped to the cell; IEEE-exact float operations only, so the CPU
 * restatement produces the same bits.  xyz: 9 * ntris floats, idx: 3 * ntris (0, 1, 2, ...) or NULL, rgb: 3 * ntris
 * packed colours (hash of the vertex index) or NULL — DEVICE pointers on the current CUDA device; enqueued on
 * cuda_stream.  Needs no spraygpu_device. */
SPRAYGPU_API int spraygpu_gen_soup(uint32_t key, size_t ntris, const float lo[3], const float hi[3], float h, float* xyz,
                                   uint32_t* idx, uint32_t* rgb, void* cuda_stream);

/* Per-vertex normals WITHOUT any allocation.  No counterpart in the reference (Embree and the reference allocate
 * freely on the host); it exists for hosts that cannot let a plug-in call cudaMalloc — typically the simulation an
 * in-situ renderer lives in, which has given all of HBM to its own memory pool (a cudaMalloc would fail or eat into
 * the solver's budget) and whose streams a cudaMalloc / cudaFree would stall with their device-wide synchronisation.
 * Same kernels and bit-identical normals as spraygpu_scene_commit + spraygpu_scene_get_normals (no transform), but
 * the work space (valences, face lists) is a fixed array inside the library image, not a grown scratch buffer:
 *   - limits: nv <= SPRAYGPU_STATIC_MAX_VERTS, nf <= SPRAYGPU_STATIC_MAX_FACES, else SPRAYGPU_ERR_LIMIT; every GPU
 *     that loads the library carries the array (about 4.6 MB) until the process exits;
 *   - xyz (nv*3 floats), idx (nf*3 uint32) and normals (nv*3 floats, overwritten) are DEVICE pointers on the current
 *     CUDA device — a host array would need a staging allocation and is refused with SPRAYGPU_ERR_ARG; needs no
 *     spraygpu_device (creating one allocates); cuda_stream is the caller's cudaStream_t (NULL: default stream);
 *   - face indices are validated like spraygpu_scene_set_mesh does: the call waits on the stream once for a 4-byte
 *     read-back, then enqueues the normals kernels and returns without waiting for them;
 *   - the first call on each GPU also looks up the array's device address; make that call (any small mesh will do)
 *     during set-up, before any stream capture begins;
 *   - all calls on a GPU share the one array: calls whose kernels could overlap (two streams, two threads) race —
 *     serialise them. */
#define SPRAYGPU_STATIC_MAX_VERTS (1u << 17)
#define SPRAYGPU_STATIC_MAX_FACES (1u << 18)
SPRAYGPU_API int spraygpu_compute_normals_static(const float* xyz, size_t nv, const uint32_t* idx, size_t nf, float* normals,
                                                 void* cuda_stream);

/* Scene::intersect on ONE domain = rtcIntersect + TriMeshBuffer::updateIntersection, batched
 * (src/render/scene.inl:190-199).  keys[i] carries domain_id in its domain field; uv may be NULL. [h|d] */
SPRAYGPU_API int spraygpu_intersect(spraygpu_scene scene, int domain_id, const spraygpu_ray* rays, size_t n, spraygpu_key* keys,
                       float* uv);
/* Scene::occluded = rtcOccluded, batched (src/render/scene.inl:202-209).  occ[i] = 1 if occluded. [h|d] */
SPRAYGPU_API int spraygpu_occluded(spraygpu_scene scene, const spraygpu_ray* rays, size_t n, uint8_t* occ);

/* ---- domain table: replaces WbvhEmbree::init/build (src/render/wbvh_embree.cc:31-111) and
 *      Scene::getBsdf (domain.h:34) ---- */
SPRAYGPU_API int spraygpu_domains_set(spraygpu_device dev, int ndomains, const float* world_aabb6, const spraygpu_material* mtl);
/* Bind a committed scene as the resident data of a domain (Scene::load cache hit, scene.inl:135-187);
 * scene == NULL marks the domain non-resident on this GPU (in-situ partition / cache eviction). */
SPRAYGPU_API int spraygpu_domain_bind(spraygpu_device dev, int domain_id, spraygpu_scene scene);

/* Ray -> list of ALL entered domains sorted by (tmin, id): WbvhEmbree::intersect + cbIntersect1 +
 * DomainList::sort (wbvh_embree.cc:126-140, rays.h:71-79, aabb.h:139-169).  cap entries per ray. [h|d] */
SPRAYGPU_API int spraygpu_domain_lists(spraygpu_device dev, const spraygpu_ray* rays, size_t n, int cap, int32_t* counts,
                          int32_t* ids, float* tmins);
/* Closest hit over all RESIDENT entered domains (what the speculative queues + VBuf converge to).
 * merge = 0: keys[i] is overwritten; merge = 1: keys[i] = min(keys[i], local result) with a 64-bit
 * atomicMin (multi-GPU / out-of-core accumulation).  uv may be NULL. [h|d] */
SPRAYGPU_API int spraygpu_intersect_domains(spraygpu_device dev, const spraygpu_ray* rays, size_t n, spraygpu_key* keys,
                               float* uv, int merge);
/* Any hit over all RESIDENT entered domains.  merge = 1: occ[i] |= local. [h|d] */
SPRAYGPU_API int spraygpu_occluded_domains(spraygpu_device dev, const spraygpu_ray* rays, size_t n, uint8_t* occ, int merge);
/* TriMeshBuffer::updateIntersection for resolved keys (trimesh_buffer.cc:328-360): packed colour and
 * unnormalised shading normal; u,v are recomputed from (ray, primID).  [h|d] */
SPRAYGPU_API int spraygpu_hit_attributes(spraygpu_device dev, const spraygpu_ray* rays, const spraygpu_key* keys, size_t n,
                            uint32_t* color, float* ns3);
/* Eye rays of a blocking tile (ooc::Tracer::genSingleEyes/genMultiEyes, ooc_tracer.inl:84-172).  [h|d] */
SPRAYGPU_API int spraygpu_gen_eye_rays(spraygpu_device dev, const spraygpu_camera* cam, spraygpu_tile tile, int spp,
                          spraygpu_ray* rays);
/* The in-situ tracers' eye rays (insitu::genSingleSampleEyeRays / genMultiSampleEyeRays,
 * src/insitu/insitu_ray.h:103-182): same layout, jitter seeded with (pixid, s).  [h|d] */
SPRAYGPU_API int spraygpu_gen_eye_rays_insitu(spraygpu_device dev, const spraygpu_camera* cam, spraygpu_tile tile, int spp,
                                              spraygpu_ray* rays);

/* ---- the fused path: one blocking tile from eye rays to radiance in the frame buffer.
 * Replaces ooc::Tracer::trace's per-tile body (ooc_tracer.inl:184-234): genSingleEyes/genMultiEyes,
 * Isector (domain culling), PContext::isectPrims (rtcIntersect/rtcOccluded loops, VBuf, shaders) and
 * TContext::retire -> HdrImage::add (ooc_tcontext.inl:123-135, display/image.h:99-109).
 * fb: image_w*image_h float4, ACCUMULATED into (caller clears, as SprayRenderer::renderFilm does). [h|d]
 * stats may be NULL. */
SPRAYGPU_API int spraygpu_trace_tile(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                        spraygpu_tile tile, float* fb_rgba, spraygpu_stats* stats);
/* Same call, but with a HOST frame buffer it returns as soon as the tile is enqueued: the tile's rows reach
 * fb after spraygpu_device_synchronize(dev).  The D2H copy of tile k runs on a second stream and overlaps the
 * kernel of tile k+1 (a film-mode frame = a loop of these + one synchronize).  fb should be pinned memory. */
SPRAYGPU_API int spraygpu_trace_tile_async(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                                           spraygpu_tile tile, float* fb_rgba);
/* Same, with BVH node / triangle fetch counters (slower; used for the roofline's N_node, N_tri). */
SPRAYGPU_API int spraygpu_trace_tile_counted(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                                spraygpu_tile tile, float* fb_rgba, spraygpu_stats* stats);

/* A whole frame = ooc::Tracer::trace's loop over the blocking tiles (ooc_tracer.inl:184-234) in ONE persistent
 * launch (per 32 tiles): the tiles' sample slots are concatenated, so there is one ramp-up / tail per frame instead of
 * one per tile; samid stays the index within its tile.  Same result as calling spraygpu_trace_tile per tile.
 * With a host fb it falls back to per-tile launches whose D2H copies overlap the next tile (and synchronises). [h|d] */
SPRAYGPU_API int spraygpu_trace_frame(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                                      const spraygpu_tile* tiles, int ntiles, float* fb_rgba, spraygpu_stats* stats);

/* The same for RECTANGLES of blocking tiles: rects[i] (inside blocking_tiles[i]) is rendered, but every sample keeps the
 * samid it has in its blocking tile, so seeds (samid * depth, ooc_shader_pt.h:130,217; bufid jitter, ooc_tracer.inl:152)
 * and therefore the image do not depend on how a frame is cut across GPUs.  (The reference's --partition image numbers
 * samples per rank-local stripe, src/render/tile.cc:317-391, so ITS multi-sample / multi-bounce noise pattern changes
 * with the rank count.)  fb must be a device pointer. */
SPRAYGPU_API int spraygpu_trace_frame_rects(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                                            const spraygpu_tile* rects, const spraygpu_tile* blocking_tiles, int n,
                                            float* fb_rgba, spraygpu_stats* stats);

/* ---- out-of-core: the device cache holds fewer domains than the tile enters.
 * Replaces ooc::PContext::isectPrims' domain loop (src/ooc/ooc_pcontext.h:92-181): per round, schedule the
 * domains by queued rays (DomainStats::schedule, src/ooc/ooc_domain_stats.cc:60-111), `scene->load(id)` each one
 * (LruCache miss -> PlyLoader + rtcCommit, :145-154) and drain its radiance + shadow queues.  Here the tile runs
 * as a wavefront swept over groups of <= max_resident domains; per-ray state is merged across groups with a
 * 64-bit atomicMin on the hit keys and an OR on the occlusion bytes.  bounces + 1 sweeps per tile.
 *
 * spraygpu_load_fn = Scene::load (src/render/scene.inl:135-187): on return 0 the domain must be bound on `dev`
 * (spraygpu_domain_bind); the callback may evict (unbind) other domains, but none it loaded since the current
 * group started — true for an LRU cache of capacity >= max_resident.  Called on the calling thread.
 * Same result as spraygpu_trace_tile with every domain resident.  fb must be a device pointer. */
typedef int (*spraygpu_load_fn)(void* user, int domain_id);
SPRAYGPU_API int spraygpu_trace_tile_ooc(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                                         spraygpu_tile tile, float* fb_rgba, int max_resident, spraygpu_load_fn load,
                                         void* user, spraygpu_stats* stats);

/* ---- multi-GPU: one process (rank) per GPU on one NVLink/NVSwitch box.  Replaces the reference's MPI
 *      (SURVEY.md section 2.2): MPI_Init/Comm_rank (src/apps/spray_insitu_*.cc:33-38), the in-situ domain
 *      partition (src/render/data_partition.h:59-137), the t-buffer MPI_Allreduce(MIN)
 *      (src/insitu/insitu_vbuf.h:113), the occlusion MPI_Allreduce(BOR) (:124), the MPI_Isend/Recv ray
 *      forwarding (src/insitu/insitu_comm.inl:40-98) and HdrImage::composite's MPI_Reduce(SUM)
 *      (src/display/image.h:167-181).  NCCL is loaded at run time (libnccl.so.2); without it these
 *      entry points return SPRAYGPU_ERR_STATE. ---- */
typedef struct spraygpu_comm_s* spraygpu_comm;
#define SPRAYGPU_UNIQUE_ID_BYTES 128
/* ncclGetUniqueId: rank 0 creates it, the host application broadcasts the 128 bytes (any transport). */
SPRAYGPU_API int spraygpu_comm_unique_id(char id[SPRAYGPU_UNIQUE_ID_BYTES]);
/* ncclCommInitRank on dev's GPU (collective over all ranks). */
SPRAYGPU_API int spraygpu_comm_create(spraygpu_device dev, int nranks, int rank, const char id[SPRAYGPU_UNIQUE_ID_BYTES],
                                      spraygpu_comm* out);
SPRAYGPU_API int spraygpu_comm_destroy(spraygpu_comm comm);
/* InsituPartition::partition (data_partition.h:59-137; morton.h:33-52): 30-bit Morton code of each
 * domain's AABB centre normalised to the scene AABB, sorted, contiguous runs of ndomains/nranks per rank;
 * round_robin != 0 = the reference's `#else` branch (:139-155).  Pure host function (no GPU needed).
 * codes may be NULL. */
SPRAYGPU_API int spraygpu_partition_domains(int ndomains, const float* world_aabb6, const float scene_aabb6[6], int nranks,
                                            int round_robin, int32_t* domain_to_rank, uint32_t* codes);
/* The in-situ blocking tile: same contract and result as spraygpu_trace_tile, but every rank holds only
 * ITS domains (spraygpu_domain_bind) and all ranks call this collectively with identical arguments.
 * Per ray generation: eye / bounce rays are identical on every rank; each rank traces them against its
 * resident domains; hits are resolved by ncclAllReduce(ncclUint64, ncclMin) on the packed keys
 * (insitu::VBuf::compositeTbuf); the rank owning the winning domain shades; shadow directions travel by
 * an all-reduce, occlusion by ncclAllReduce(max) on one byte per ray (compositeObuf); the owner adds the
 * radiance into ITS frame buffer.  fb therefore holds this rank's partial image: sum the ranks with
 * spraygpu_image_reduce.  fb must be a device pointer.  stats (may be NULL) are summed over ranks with a collective:
 * pass NULL on every rank or non-NULL on every rank (the same holds for the _p2p and _fwd variants below).
 * insitu::{SingleThread,MultiThread}Tracer::trace (src/insitu/insitu_singlethread_tracer.inl:381-466). */
SPRAYGPU_API int spraygpu_trace_tile_insitu(spraygpu_device dev, spraygpu_comm comm, const spraygpu_camera* cam,
                                            const spraygpu_shade* shade, spraygpu_tile tile, float* fb_rgba,
                                            spraygpu_stats* stats);
/* The same blocking tile with the exchange done over PEER MEMORY (CUDA IPC mappings of every rank's wave buffers over
 * NVLink / NVSwitch) inside the compute kernels instead of dense NCCL all-reduces: a hit merges its packed key into
 * every rank's key array with a 64-bit system-scope atomicMin, the rank owning the winning domain stores 16 B of hit
 * attributes to every rank, every rank shades every hit (identical arithmetic, nothing else travels), occlusions are
 * single-byte stores to the sample's HOME rank, and ranks meet at device-side barriers (no host involvement per wave).
 * The home rank of image row y is (y / rows_per_block) % nranks: it alone accumulates that row, so fb ends up holding
 * this rank's ROWS of the final image (not a partial sum) — describe the same blocks to spraygpu_image_set_rows and
 * spraygpu_image_flush moves them to the root / host.  Same image, statistics and sample numbering as
 * spraygpu_trace_tile_insitu.  One rank per GPU; every GPU must be able to map its peers (NVLink or PCIe P2P). */
SPRAYGPU_API int spraygpu_trace_tile_insitu_p2p(spraygpu_device dev, spraygpu_comm comm, const spraygpu_camera* cam,
                                                const spraygpu_shade* shade, spraygpu_tile tile, float* fb_rgba,
                                                int rows_per_block, spraygpu_stats* stats);
/* The in-situ blocking tile with TRUE RAY FORWARDING — replaces insitu::MultiThreadTracer's send / receive loop
 * (src/insitu/insitu_multithread_tracer.inl:135-195, :313-442), insitu::Comm's per-destination MsgHeader + Ray[]
 * messages (src/insitu/insitu_comm.inl:51-92) and the WorkStats termination count (src/insitu/insitu_work_stats.cc:34-85).
 * Every sample has a HOME rank (image row y -> (y / rows_per_block) % nranks, as in spraygpu_trace_tile_insitu_p2p): the
 * home rank alone generates, shades and accumulates it.  A ray is appended — 32 bytes, from inside the kernel that
 * produced it, over NVLink / NVSwitch peer memory — to the inbox of every rank that owns a domain the ray ENTERS (exact
 * slab test against all domain boxes); owners trace their inbox against their own domains, merge closest hits into the
 * home rank's key array with a 64-bit system-scope atomicMin, store occlusions as single bytes and hit attributes as
 * 16-byte records to the home rank.  Device-side barriers only; the number of waves is `bounces` (+ 1 for the trailing
 * shadow rays).  domain_owner[i] = rank that holds domain i (spraygpu_partition_domains); a rank must have exactly its
 * own domains resident.  fb receives this rank's HOME rows (describe them to spraygpu_image_set_rows, flush with
 * spraygpu_image_flush).  Same image, statistics and sample numbering as spraygpu_trace_tile_insitu(_p2p). */
SPRAYGPU_API int spraygpu_trace_tile_insitu_fwd(spraygpu_device dev, spraygpu_comm comm, const spraygpu_camera* cam,
                                                const spraygpu_shade* shade, spraygpu_tile tile, float* fb_rgba,
                                                int rows_per_block, const int32_t* domain_owner /* ndomains */,
                                                spraygpu_stats* stats);
/* HdrImage::composite (image.h:167-181): fb = sum over ranks, valid on `root` (in place, device pointer). */
SPRAYGPU_API int spraygpu_image_reduce(spraygpu_device dev, spraygpu_comm comm, float* fb_rgba, size_t npixels, int root);
/* Image-partition composite for DISJOINT row bands (rank r rendered rows [row_begin[r], row_begin[r+1])):
 * root receives every band in place (ncclSend/ncclRecv group) — 1/nranks of the reduce's traffic. */
SPRAYGPU_API int spraygpu_image_gather_rows(spraygpu_device dev, spraygpu_comm comm, float* fb_rgba, int image_w,
                                            const int32_t* row_begin /* nranks + 1 */, int root);
/* The same for INTERLEAVED row blocks (block b = rows [block_row_begin[b], block_row_begin[b+1]) rendered by rank
 * block_owner[b]): image-space load balancing — a contiguous band per rank leaves the ranks that own sky idle. */
SPRAYGPU_API int spraygpu_image_gather_blocks(spraygpu_device dev, spraygpu_comm comm, float* fb_rgba, int image_w, int nblocks,
                                              const int32_t* block_row_begin /* nblocks + 1 */,
                                              const int32_t* block_owner /* nblocks */, int root);
/* ---- the frame image: replaces HdrImage (src/display/image.h:58-239: buf, clear(), composite() = MPI_Reduce(SUM) of
 *      every rank's whole image into rank 0, :167-181) on the GPU path.  A frame is traced into a DEVICE render target;
 *      spraygpu_image_flush then moves the rows THIS rank rendered — one kernel of this library, 16-byte stores on
 *      peer-mapped (CUDA IPC over NVLink) / host-mapped (pinned, PCIe) pointers, no collective call per frame — and
 *      zeroes the render target (the next frame's image.clear()).  Only pixels that are non-zero, or were non-zero
 *      when last flushed to that destination, are written; the destination is bit-identical to a full copy as long as
 *      nothing else writes it.
 *      comm == NULL: single GPU (no root composite).  shm_name != NULL: the host image is a POSIX shared-memory
 *      segment registered with CUDA in every rank (all ranks of the box write ONE host image, each over its own PCIe
 *      link); NULL: a private pinned allocation.  image_w must be a multiple of 32.  Collective when comm != NULL. */
typedef struct spraygpu_image_s* spraygpu_image;
SPRAYGPU_API int spraygpu_image_create(spraygpu_device dev, spraygpu_comm comm, int image_w, int image_h, const char* shm_name,
                                       spraygpu_image* out);
SPRAYGPU_API int spraygpu_image_destroy(spraygpu_image img);                 /* collective when created with a comm */
SPRAYGPU_API float* spraygpu_image_device(spraygpu_image img);               /* this rank's render target (device, w*h float4) */
SPRAYGPU_API float* spraygpu_image_host(spraygpu_image img);                 /* HdrImage::buf: pinned host pixels */
SPRAYGPU_API float* spraygpu_image_root(spraygpu_image img);                 /* the root's composite as mapped in THIS process (device) */
/* Which rows this rank renders: blocks [row_begin[b], row_begin[b+1]) with owner[b] == rank (owner NULL: all mine;
 * owner[b] == -1: rendered by nobody on this communicator).
 * Replaces the rank's stripe of ImageScheduleTileList (src/render/tile.cc:317-391).
 * May be called again later (re-balancing): rows this rank gives up are RELEASED — what its earlier flushes left
 * non-zero in them at the root composite / the host image is zeroed there, after pending asynchronous flushes have
 * landed — so the new owner starts from black.  Call it between a flush and the next trace (unflushed pixels of released
 * rows stay in the render target).  Collective when the image was created with a comm (every rank calls it; ranks
 * meet before anyone flushes under the new ownership). */
SPRAYGPU_API int spraygpu_image_set_rows(spraygpu_image img, int nblocks, const int32_t* row_begin /* nblocks + 1 */,
                                         const int32_t* owner /* nblocks or NULL */);
/* HdrImage::composite (to_root) and/or the download into HdrImage::buf (to_host) for this rank's rows, then
 * image.clear() of the render target (clear).  Enqueued on the device's stream. */
SPRAYGPU_API int spraygpu_image_flush(spraygpu_image img, int to_root, int to_host, int clear);
/* The same flush (target always cleared), with the PCIe half OVERLAPPED with the next frame: on the device's stream the
 * render target is scanned, root-bound pixels are pushed over NVLink, host-bound pixels are compacted into a device list;
 * a few small CTAs on the image's own stream then stream that list into the pinned host image while the caller's next
 * spraygpu_trace_frame runs.  spraygpu_image_wait: host wait until every flush issued so far has landed (renderFilm's
 * loop, src/render/spray_renderer.inl:355-372, with composite() of frame k hidden behind trace() of frame k + 1). */
SPRAYGPU_API int spraygpu_image_flush_async(spraygpu_image img, int to_root, int to_host);
SPRAYGPU_API int spraygpu_image_wait(spraygpu_image img);
/* Device-side join (no host wait): work enqueued on the device's stream afterwards starts when every flush issued so far has landed. */
SPRAYGPU_API int spraygpu_image_join(spraygpu_image img);
/* cumulative pixels (16 B each) stored to the root composite / the host image by this rank's flushes */
SPRAYGPU_API int spraygpu_image_flushed_pixels(spraygpu_image img, uint64_t* to_root, uint64_t* to_host);
/* Plain full-size D2H copy of this rank's render target (which = 0) or of the root composite as mapped here
 * (which = 1) into out_host (w*h float4); synchronises.  For checks and PPM output, not the frame path. */
SPRAYGPU_API int spraygpu_image_read(spraygpu_image img, int which, float* out_host);
/* stream synchronise + rendezvous of all ranks: after it every rank's flushed rows are visible at their destination */
SPRAYGPU_API int spraygpu_image_barrier(spraygpu_image img);

/* ncclAllReduce(ncclUint64, ncclMin) on hit keys / ncclAllReduce(ncclUint8, ncclMax) on occlusion bytes,
 * in place on device buffers: the building blocks for callers that drive spraygpu_intersect_domains /
 * spraygpu_occluded_domains themselves. */
SPRAYGPU_API int spraygpu_allreduce_keys_min(spraygpu_device dev, spraygpu_comm comm, spraygpu_key* keys, size_t n);
SPRAYGPU_API int spraygpu_allreduce_occ_max(spraygpu_device dev, spraygpu_comm comm, uint8_t* occ, size_t n);

#ifdef __cplusplus
}
#endif
#endif
