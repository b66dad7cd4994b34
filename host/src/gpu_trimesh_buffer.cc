// gpu_trimesh_buffer.cc — see spray/gpu_trimesh_buffer.h.
#include "spray/gpu_trimesh_buffer.h"

#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "spray/check.h"
#include "spray/render_types.h"

namespace spray {

void GpuTriMeshBuffer::cleanup() {
  for (spraygpu_scene s : scenes_)
    if (s) spraygpu_scene_destroy(s);
  scenes_.clear();
  freePinned(vertices_); freePinned(faces_); freePinned(colors_);
  vertices_ = nullptr; faces_ = nullptr; colors_ = nullptr;
}

void GpuTriMeshBuffer::init(int max_cache_size_ndomains, std::size_t max_nvertices, std::size_t max_nfaces,
                            bool compute_normals) {
  SPRAY_CHECK(device_ != nullptr, "GpuTriMeshBuffer::setDevice must be called before init");
  SPRAY_CHECK(max_cache_size_ndomains > 0, "");
  SPRAY_CHECK(max_nfaces <= SPRAYGPU_MAX_PRIMS, "domain with " << max_nfaces << " faces exceeds the 23-bit primID field");
  cleanup();
  max_cache_size_ = max_cache_size_ndomains;
  max_nvertices_ = max_nvertices;
  max_nfaces_ = max_nfaces;
  compute_normals_ = compute_normals;  // the GPU commit always produces normals (shading needs them)
  const std::size_t n = (std::size_t)max_cache_size_ndomains;
  num_vertices_.assign(n, 0);
  num_faces_.assign(n, 0);
  block_domain_.assign(n, -1);
  // ONE pinned staging block instead of cache_size x max-sized host arrays (trimesh_buffer.cc:70-92): the
  // geometry's home is HBM, the host copy is only the PLY decode target
  vertices_ = (float*)allocPinned(sizeof(float) * 3 * max_nvertices);
  faces_ = (uint32_t*)allocPinned(sizeof(uint32_t) * 3 * max_nfaces);
  colors_ = (uint32_t*)allocPinned(sizeof(uint32_t) * max_nvertices);
  SPRAY_CHECK(vertices_ && faces_ && colors_, "staging allocation failed");
  scenes_.assign(n, nullptr);
  for (std::size_t i = 0; i < n; ++i) SPRAY_GPU_CALL(spraygpu_scene_create(device_, &scenes_[i]));
}

spraygpu_scene GpuTriMeshBuffer::load(const std::string& filename, int cache_block, const mat4& transform,
                                      bool apply_transform) {
  SPRAY_CHECK(cache_block >= 0 && cache_block < max_cache_size_, "cache block " << cache_block);
  const auto t0 = std::chrono::steady_clock::now();
  PlyLoader::Data d;
  d.vertices_capacity = max_nvertices_ * 3;
  d.faces_capacity = max_nfaces_ * 3;
  d.colors_capacity = max_nvertices_;
  d.vertices = vertices_;
  d.faces = faces_;
  d.colors = colors_;
  d.num_vertices = d.num_faces = 0;
  // a PLY without colours leaves the reference's colour array untouched (stale arena memory); zero here
  std::memset(colors_, 0, sizeof(uint32_t) * max_nvertices_);
  loader_.load(filename, &d);
  const auto t1 = std::chrono::steady_clock::now();
  num_vertices_[(std::size_t)cache_block] = d.num_vertices;
  num_faces_[(std::size_t)cache_block] = d.num_faces;

  spraygpu_scene sc = scenes_[(std::size_t)cache_block];
  SPRAY_GPU_CALL(spraygpu_scene_set_mesh(sc, d.vertices, d.num_vertices, d.faces, d.num_faces, d.colors));
  SPRAY_GPU_CALL(spraygpu_scene_set_transform(sc, apply_transform ? transform.data() : nullptr));
  SPRAY_GPU_CALL(spraygpu_scene_commit(sc));
  const auto t2 = std::chrono::steady_clock::now();
  last_parse_ms_ = std::chrono::duration<double, std::milli>(t1 - t0).count();
  last_build_ms_ = std::chrono::duration<double, std::milli>(t2 - t1).count();
  return sc;
}

void GpuTriMeshBuffer::updateIntersection(int cache_block, RTCRayIntersection* isect) const {
  const int dom = block_domain_[(std::size_t)cache_block];
  SPRAY_CHECK(dom >= 0, "cache block " << cache_block << " is not bound to a domain");
  spraygpu_ray r = {isect->org[0], isect->org[1], isect->org[2], isect->tnear,
                    isect->dir[0], isect->dir[1], isect->dir[2], INFINITY};
  uint32_t tbits;
  std::memcpy(&tbits, &isect->tfar, 4);
  const spraygpu_key key = ((uint64_t)tbits << 32) | ((uint64_t)(unsigned)dom << 23) | (uint64_t)isect->primID;
  SPRAY_GPU_CALL(spraygpu_hit_attributes(device_, &r, &key, 1, &isect->color, isect->Ns));
}

}  // namespace spray
