// internal.cuh — host-side state of libspraygpu (device context, per-domain scene, domain table).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "../../include/spray_gpu.h"
#include "core.cuh"

namespace spraygpu {

void set_error(const char* fmt, ...);

#define SG_CUDA(call)                                                                          \
  do {                                                                                         \
    cudaError_t e__ = (call);                                                                  \
    if (e__ != cudaSuccess) {                                                                  \
      set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));   \
      return SPRAYGPU_ERR_CUDA;                                                                \
    }                                                                                          \
  } while (0)

#define SG_CHECK(cond, code, ...)            \
  do {                                       \
    if (!(cond)) {                           \
      set_error(__VA_ARGS__);                \
      return code;                           \
    }                                        \
  } while (0)

// Kernels that stage the domain table + top-level tree in dynamic shared memory need the opt-in above 48 KB
// (hundreds of domains with a deep top-level tree); refuse what no B200 CTA can hold.
#define SG_DYN_SMEM(kern, bytes)                                                                                     \
  do {                                                                                                               \
    const size_t b__ = (bytes);                                                                                      \
    SG_CHECK(b__ <= 227u * 1024u, SPRAYGPU_ERR_LIMIT, "%s: %zu B of shared memory for the domain table (limit 227 KB)", #kern, b__); \
    if (b__ > 48u * 1024u) SG_CUDA(cudaFuncSetAttribute((const void*)(kern), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b__)); \
  } while (0)

// device-visible per-domain record (96 B)
struct DomainEntry {
  float box[6];           // world AABB as given by the scene description
  int32_t resident;       // 1 if a committed scene is bound on THIS GPU
  int32_t mtl_type;       // SPRAYGPU_MTL_*
  float mtl[3];
  uint32_t nprims;
  const uint4* nodes;     // BVH8 nodes (5 x uint4 each)
  const float4* tris;     // 3 x float4 per triangle, BVH leaf order
  const float4* attr;     // 3 x float4 per triangle, primID order: (n_k.xyz, rgb_k bits)
  const float* verts;     // world-space vertices (key -> (u,v) re-derivation)
  const uint32_t* faces;  // index triples, primID order
  uint64_t pad;
};
static_assert(sizeof(DomainEntry) == 96, "DomainEntry layout");

// Top-level BVH4 over the world boxes of the RESIDENT domains (rebuilt on the host whenever the table
// changes; <= 512 leaves).  Plain float boxes, SoA per axis so a node is 7 x 16-byte shared-memory loads.
// child >= 0: inner node index; child < 0: ~domain id; unused slots carry an inverted box.
struct TlasNode {
  float lo[3][4], hi[3][4];
  int32_t child[4];
};
static_assert(sizeof(TlasNode) == 112, "TlasNode layout");
#define SPRAY_TLAS_MAX_NODES 384

struct Scratch {
  void* ptr = nullptr;
  size_t bytes = 0;
};

struct Scene;

struct Device {
  int cuda_device = 0;
  cudaStream_t stream = nullptr;   // stream every launch goes to (default: legacy default stream)
  bool own_stream = false;
  int sm_count = 148;
  Scratch scratch;                 // build scratch (grown on demand)
  Scratch stage_in, stage_out;     // device staging for host-pointer calls
  Scratch wave;                    // in-situ wavefront buffers (insitu.cu)
  // host frame buffers: two device staging images, D2H on a copy stream so that tile k's copy overlaps
  // tile k+1's kernel
  Scratch fb_stage[2];
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_render[2] = {nullptr, nullptr}, ev_copy[2] = {nullptr, nullptr};
  uint64_t fb_calls = 0;
  void* pinned = nullptr;          // pinned bounce buffer for small D2H reads
  size_t pinned_bytes = 0;
  // domain table
  int ndomains = 0;
  std::vector<DomainEntry> h_domains;
  std::vector<Scene*> bound;
  DomainEntry* d_domains = nullptr;
  bool table_dirty = true;
  // out-of-core sweeps (ooc.cu): when non-empty, only resident domains with active[id] != 0 are visible to
  // the kernels (uploaded table + top-level BVH); empty = every resident domain
  std::vector<uint8_t> active;
  std::vector<TlasNode> h_tlas;    // built in upload_domain_table
  TlasNode* d_tlas = nullptr;
  int tlas_n = 0;                  // 0: no resident domain
  // acceleration-structure builds and mesh uploads go to `load_stream` while `loading_async` is set (the out-of-core
  // tile prefetches the next domain group while the current one is traced, ooc.cu); otherwise to `stream`
  cudaStream_t load_stream = nullptr;     // builds of prefetched domains
  cudaStream_t upload_stream = nullptr;   // their mesh uploads: domain k + 1 crosses PCIe while domain k builds
  bool loading_async = false;
  std::vector<uint8_t> tracing;           // domains of the group being traced while the next one loads: must not be evicted
  int collapse_blocks_per_sm = 0;  // occupancy of the cooperative collapse kernel (queried once)
  // small domains (<= SPRAY_POOL_MAX_TRIS triangles) build on a pool of streams, each with its own scratch: a 5 480-triangle
  // commit is ~30 launches of near-empty kernels (0.25 ms of latency, the GPU idle), so enqueue-only commits of many small
  // domains overlap each other instead of queueing on one stream (C2 / C5 scene load: 64 wavelet domains)
  cudaStream_t pool_stream[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  Scratch pool_scratch[8];
  cudaEvent_t pool_order = nullptr;
  unsigned pool_rr = 0;
  unsigned long long* d_counters = nullptr;  // [0] work counter, [1..] stats
  uint64_t launches = 0;           // kernels launched by this library (bench "gpu_launches")
  // last render-kernel configuration (function attribute + occupancy are queried once per kernel / smem size)
  const void* render_kern = nullptr;
  size_t render_smem = 0;
  int render_per_sm = 0;
};

struct Scene {
  Device* dev = nullptr;
  size_t nv = 0, nf = 0;
  float* d_verts_obj = nullptr;  // as given by set_mesh
  uint32_t* d_faces = nullptr;
  uint32_t* d_rgb = nullptr;
  bool has_xf = false;
  float xf[16];
  // committed
  bool committed = false;
  float* d_verts = nullptr;    // world space (aliases d_verts_obj when no transform, else d_verts_xf)
  float* d_verts_xf = nullptr;
  // capacities (bytes) of the buffers above: a cache block that is re-loaded with a domain of the same size (the
  // out-of-core LRU path) re-uses its allocations instead of cudaFree (a device-wide sync) + cudaMalloc
  size_t cap_verts_obj = 0, cap_faces = 0, cap_rgb = 0, cap_verts_xf = 0, cap_normals = 0, cap_tris = 0, cap_attr = 0,
         cap_nodes = 0;
  float* d_normals = nullptr;  // nv*3
  uint4* d_nodes = nullptr;
  uint32_t n_nodes = 0;
  uint32_t bvh_depth = 0;      // BVH8 levels (<= SPRAY_STACK_SIZE, checked at commit)
  float4* d_tris = nullptr;
  float4* d_attr = nullptr;
  float bounds[6];
  double build_ms = 0.0;
  // an enqueued build whose results have not been read back yet (build_scene / scene_finalize)
  bool pending = false;
  uint32_t node_cap = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t ev_upload = nullptr; // set_mesh's copies (recorded on the stream they went to)
  uint32_t* h_result = nullptr;    // pinned: 8 counter words + 12 bounds words
  // CUDA graph of the build's launches for the shape `graph_key` (captured on the second commit of that shape)
  cudaGraphExec_t graph_exec = nullptr;
  uint64_t graph_key = 0, seen_key = 0;
  uint32_t graph_launches = 0;
  bool graph_failed = false;
  int pool_slot = -1;              // build-stream pool slot of a small domain (Device::pool_stream)
};
#define SPRAY_BUILD_POOL 8
#define SPRAY_POOL_MAX_TRIS (1u << 16)
inline cudaStream_t build_stream(const Device* d) { return d->loading_async && d->load_stream ? d->load_stream : d->stream; }
inline cudaStream_t upload_stream(const Device* d) { return d->loading_async && d->upload_stream ? d->upload_stream : d->stream; }

int ensure_scratch(Scratch& s, size_t bytes);
// make *p hold at least `bytes` (re-using the allocation when it is large enough and not more than 2x too large)
template <typename T>
inline int grow_buffer(T** p, size_t* cap, size_t bytes) {
  if (*p && *cap >= bytes && *cap <= 2 * bytes + 4096) return SPRAYGPU_OK;
  if (*p) cudaFree(*p);
  *p = nullptr; *cap = 0;
  SG_CUDA(cudaMalloc((void**)p, bytes));
  *cap = bytes;
  return SPRAYGPU_OK;
}
int build_scene(Scene* sc);     // build.cu: enqueue the build (no host wait)
int scene_finalize(Scene* sc);  // build.cu: wait for it, read node count / depth / bounds, run the commit's checks
int launch_max_index(Device* dev, const uint32_t* d_idx, size_t n, uint32_t* d_out);   // build.cu
int upload_domain_table(Device* dev);
CameraDev to_camdev(const spraygpu_camera* c);   // trace.cu

// trace.cu launchers
int launch_domain_lists(Device* dev, const spraygpu_ray* rays, size_t n, int cap, int32_t* counts, int32_t* ids,
                        float* tmins);
int launch_intersect_scene(Scene* sc, int domain_id, const spraygpu_ray* rays, size_t n, uint64_t* keys, float* uv);
int launch_occluded_scene(Scene* sc, const spraygpu_ray* rays, size_t n, uint8_t* occ);
int launch_intersect_domains(Device* dev, const spraygpu_ray* rays, size_t n, uint64_t* keys, float* uv, int merge);
int launch_occluded_domains(Device* dev, const spraygpu_ray* rays, size_t n, uint8_t* occ, int merge);
int launch_hit_attributes(Device* dev, const spraygpu_ray* rays, const uint64_t* keys, size_t n, uint32_t* color,
                          float* ns3);
int launch_eye_rays(Device* dev, const spraygpu_camera* cam, spraygpu_tile tile, int spp, int eye_seed, spraygpu_ray* rays);
int launch_render_tile(Device* dev, const spraygpu_camera* cam, const spraygpu_shade* shade, spraygpu_tile tile,
                       float* fb, bool counted);
int launch_render_tiles(Device* dev, const spraygpu_camera* cam, const spraygpu_shade* shade, const spraygpu_tile* tiles,
                        const spraygpu_tile* parents, int ntiles, float* fb, bool counted, bool reset_counters);

}  // namespace spraygpu
