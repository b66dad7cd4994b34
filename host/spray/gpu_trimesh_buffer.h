// gpu_trimesh_buffer.h — GpuTriMeshBuffer: the `SurfaceBufT` slot of Scene<CacheT, SurfaceBufT>
// (reference src/render/scene.h:62-63) filled by the B200 library instead of TriMeshBuffer + Embree
// (src/render/trimesh_buffer.h:37-108, trimesh_buffer.cc:33-360).
//
// Same four members Scene calls — init / load / get / updateIntersection — with the reference's argument
// meaning.  The "RTCScene" handed back is a spraygpu_scene (device-resident BVH8 of one cache block).
//   reference                                  here
//   rtcNewDevice                (.cc:104)      spraygpu_device (shared, set with setDevice)
//   rtcDeviceNewScene x cache   (.cc:112)      spraygpu_scene_create per cache block
//   PlyLoader::load             (.cc:133)      PlyLoader::load into pinned staging
//   x * vec4(v,1) loop          (.cc:141-157)  spraygpu_scene_set_transform (applied on the GPU at commit)
//   computeNormals              (.cc:267-326)  inside spraygpu_scene_commit (bit-exact, same summation order)
//   rtcNewTriangleMesh/rtcSetBuffer2/rtcCommit (.cc:189-225)   spraygpu_scene_set_mesh + spraygpu_scene_commit
//   updateIntersection          (.cc:328-360)  spraygpu_hit_attributes (batched) / per-hit convenience below
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "spray/ply_loader.h"
#include "spray/vecmath.h"
#include "spray_gpu.h"

namespace spray {

// The fields of RTCRayIntersection SpRay reads and writes (src/render/rays.h:246-274)
struct RTCRayIntersection {
  float org[3];
  float dir[3];
  float tnear, tfar;
  float u, v;
  unsigned geomID, primID;  // geomID: 0 on a hit, RTC_INVALID_GEOMETRY_ID on a miss
  uint32_t color;           // packed rgb, filled by updateIntersection
  float Ns[3];              // unnormalised shading normal, filled by updateIntersection
};
#define RTC_INVALID_GEOMETRY_ID ((unsigned)-1)

class GpuTriMeshBuffer {
 public:
  GpuTriMeshBuffer() {}
  ~GpuTriMeshBuffer() { cleanup(); }
  void release() { cleanup(); }  // destroy the device scenes now (before the device goes away)
  GpuTriMeshBuffer(const GpuTriMeshBuffer&) = delete;
  GpuTriMeshBuffer& operator=(const GpuTriMeshBuffer&) = delete;

  // must be called before init(): the device every cache block lives on (not owned)
  void setDevice(spraygpu_device dev) { device_ = dev; }

  void init(int max_cache_size_ndomains, std::size_t max_nvertices, std::size_t max_nfaces, bool compute_normals);

  spraygpu_scene load(const std::string& filename, int cache_block, const mat4& transform, bool apply_transform);
  spraygpu_scene get(int cache_block) { return scenes_[(std::size_t)cache_block]; }

  // colour + shading normal of ONE resolved hit (primID, u, v already in *isect); the hot path uses the
  // batched spraygpu_hit_attributes / the fused kernel instead
  void updateIntersection(int cache_block, RTCRayIntersection* isect) const;

  // extension: the domain currently held by a cache block (Scene::load records it; updateIntersection needs
  // it to address the device's domain table)
  void setBlockDomain(int cache_block, int domain_id) { block_domain_[(std::size_t)cache_block] = domain_id; }
  int blockDomain(int cache_block) const { return block_domain_[(std::size_t)cache_block]; }

  std::size_t numVertices(int cache_block) const { return num_vertices_[(std::size_t)cache_block]; }
  std::size_t numFaces(int cache_block) const { return num_faces_[(std::size_t)cache_block]; }
  double lastLoadParseMs() const { return last_parse_ms_; }
  double lastLoadBuildMs() const { return last_build_ms_; }

 private:
  void cleanup();

  spraygpu_device device_ = nullptr;
  int max_cache_size_ = 0;
  std::size_t max_nvertices_ = 0, max_nfaces_ = 0;
  bool compute_normals_ = true;
  // pinned staging for one domain (reused by every load)
  float* vertices_ = nullptr;
  uint32_t* faces_ = nullptr;
  uint32_t* colors_ = nullptr;
  std::vector<std::size_t> num_vertices_, num_faces_;
  std::vector<spraygpu_scene> scenes_;
  std::vector<int> block_domain_;
  PlyLoader loader_;
  double last_parse_ms_ = 0.0, last_build_ms_ = 0.0;
};

}  // namespace spray
