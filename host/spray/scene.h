// scene.h — Scene<CacheT, SurfaceBufT = GpuTriMeshBuffer>: the reference's scene container
// (src/render/scene.h:62-254, scene.inl:30-349) with the Embree pieces replaced by libspraygpu handles.
// Same public members and argument meaning; what the reference does per ray through rtcIntersect /
// rtcOccluded is offered batched (arrays of rays) next to the per-ray forms.
//   SceneLoader::load + mergeDomainBounds     scene.inl:37-43,296-349
//   InsituPartition::partition                 scene.inl:47-66      -> spraygpu_partition_domains
//   cache_.init / surface_buf_.init / warm-up  scene.inl:78-98
//   wbvh_.init / buildWbvh                     scene.inl:100,106-133 -> spraygpu_domains_set
//   load(id) hit / miss                        scene.inl:135-187    -> + spraygpu_domain_bind
//   intersect / occluded                       scene.inl:190-209    -> spraygpu_intersect / spraygpu_occluded
#pragma once
#include <cmath>
#include <cstring>
#include <iostream>
#include <list>
#include <string>
#include <vector>

#include "spray/caches.h"
#include "spray/check.h"
#include "spray/comm.h"
#include "spray/config.h"
#include "spray/gpu_trimesh_buffer.h"
#include "spray/scene_loader.h"
#include "spray/vecmath.h"
#include "spray_gpu.h"

namespace spray {

#define SPRAY_RAY_EPSILON 0.001f  // src/render/spray.h:46

// domain -> rank map of the in-situ mode (src/render/data_partition.h:35-160)
class InsituPartition {
 public:
  void partition(int ndomains, const std::vector<Domain>& domains, const Aabb& scene_aabb, int num_partitions) {
    std::vector<float> boxes((std::size_t)ndomains * 6);
    for (int i = 0; i < ndomains; ++i)
      for (int k = 0; k < 3; ++k) {
        boxes[6 * (std::size_t)i + k] = domains[(std::size_t)i].world_aabb.bounds[0][k];
        boxes[6 * (std::size_t)i + 3 + k] = domains[(std::size_t)i].world_aabb.bounds[1][k];
      }
    float sb[6];
    for (int k = 0; k < 3; ++k) { sb[k] = scene_aabb.bounds[0][k]; sb[3 + k] = scene_aabb.bounds[1][k]; }
    domain_to_rank_.assign((std::size_t)ndomains, 0);
    SPRAY_GPU_CALL(spraygpu_partition_domains(ndomains, boxes.data(), sb, num_partitions, 0, domain_to_rank_.data(), nullptr));
    rank_to_domains_.assign((std::size_t)num_partitions, std::list<int>());
    for (int i = 0; i < ndomains; ++i) rank_to_domains_[(std::size_t)domain_to_rank_[(std::size_t)i]].push_back(i);
  }
  int rank(int domain_id) const { return domain_to_rank_[(std::size_t)domain_id]; }
  const std::list<int>& getDomains(int rank) const { return rank_to_domains_[(std::size_t)rank]; }
  int getNumDomains(int rank) const { return (int)rank_to_domains_[(std::size_t)rank].size(); }
  const std::vector<int32_t>& computePartition() const { return domain_to_rank_; }

 private:
  std::vector<int32_t> domain_to_rank_;
  std::vector<std::list<int>> rank_to_domains_;
};

struct SceneInfo {
  spraygpu_scene rtc_scene;
  int cache_block;
};

template <typename CacheT, typename SurfaceBufT = GpuTriMeshBuffer>
class Scene {
 public:
  Scene() {}
  ~Scene() {
    for (Light* l : lights_) delete l;
    surface_buf_.release();
    if (owns_device_ && device_) spraygpu_device_destroy(device_);
  }
  Scene(const Scene&) = delete;
  Scene& operator=(const Scene&) = delete;

  // extension: the GPU this scene lives on.  Without it init() creates device `mpi::localRank()`.
  void setDevice(spraygpu_device dev) { device_ = dev; owns_device_ = false; }
  spraygpu_device device() const { return device_; }

  void init(const std::string& desc_filename, const std::string& ply_path, const std::string& storage_basepath,
            int cache_size, int view_mode, bool insitu_mode, int num_partitions) {
    SceneLoader loader;
    loader.load(desc_filename, ply_path, &domains_, &lights_);
    std::size_t max_nv, max_nf;
    mergeDomainBounds(&max_nv, &max_nf);
    SPRAY_CHECK(domains_.size() <= SPRAYGPU_MAX_DOMAINS, domains_.size() << " domains exceed the 9-bit domain field");

    insitu_ = false;
    if (view_mode == VIEW_MODE_PARTITION) {
      partition_.partition((int)getNumDomains(), getDomains(), getBound(), num_partitions);
    } else if (insitu_mode) {
      insitu_ = true;
      partition_.partition((int)getNumDomains(), getDomains(), getBound(), mpi::size());
      cache_size = partition_.getNumDomains(mpi::rank());  // scene.inl:65 "override cache size"
    }
    SPRAY_CHECK(storage_basepath.empty(), "--local-disk staging is not needed here: domains are decoded once into HBM");

    if (!device_) {
      SPRAY_GPU_CALL(spraygpu_device_create(mpi::localRank(), &device_));
      owns_device_ = true;
    }
    // the world-space domain table = WbvhEmbree::init (scene.inl:100, wbvh_embree.cc:31-56)
    std::vector<float> boxes(domains_.size() * 6);
    std::vector<spraygpu_material> mtl(domains_.size());
    for (std::size_t i = 0; i < domains_.size(); ++i) {
      for (int k = 0; k < 3; ++k) {
        boxes[6 * i + k] = domains_[i].world_aabb.bounds[0][k];
        boxes[6 * i + 3 + k] = domains_[i].world_aabb.bounds[1][k];
      }
      mtl[i].type = domains_[i].bsdf.type;
      for (int k = 0; k < 3; ++k) mtl[i].params[k] = domains_[i].bsdf.params[k];
    }
    SPRAY_GPU_CALL(spraygpu_domains_set(device_, (int)domains_.size(), boxes.data(), mtl.data()));

    if (!(view_mode == VIEW_MODE_DOMAIN || view_mode == VIEW_MODE_PARTITION)) {
      cache_.init((int)domains_.size(), cache_size, insitu_mode);
      surface_buf_.setDevice(device_);
      surface_buf_.init(cache_.getCacheSize(), max_nv, max_nf, true /* compute_normals */);
      block_domain_.assign((std::size_t)cache_.getCacheSize(), -1);
      // warm up cache (scene.inl:88-97)
      if (view_mode == VIEW_MODE_FILM || view_mode == VIEW_MODE_GLFW) {
        if (insitu_mode) {
          for (int id : partition_.getDomains(mpi::rank())) load(id);
        } else if (cache_size < 0) {
          for (std::size_t id = 0; id < domains_.size(); ++id) load((int)id);
        }
      }
    }
  }

  const InsituPartition& getInsituPartition() const { return partition_; }
  bool insitu() const { return insitu_; }

  // the domain table was uploaded in init(); kept for interface parity (SprayRenderer::init calls it)
  void buildWbvh() {}

  Aabb getBound() const {
    SPRAY_CHECK(bound_.isValid(), "invalid scene bound");
    return bound_;
  }

  // Scene::load (scene.inl:135-187): make domain `id` resident, remember it as the current domain
  void load(int id) {
    SceneInfo s;
    load(id, &s);
  }
  void load(int id, SceneInfo* sinfo) {
    int cache_block;
    if (cache_.load(id, &cache_block)) {
      scene_ = surface_buf_.get(cache_block);
      ++cache_hits_;
    } else {
      ++cache_misses_;
      const int old = block_domain_[(std::size_t)cache_block];
      if (old >= 0) SPRAY_GPU_CALL(spraygpu_domain_bind(device_, old, nullptr));  // evicted
      const mat4& x = domains_[(std::size_t)id].transform;
      const bool apply_transform = !x.isIdentity();
      scene_ = surface_buf_.load(domains_[(std::size_t)id].filename, cache_block, x, apply_transform);
      block_domain_[(std::size_t)cache_block] = id;
      surface_buf_.setBlockDomain(cache_block, id);
      SPRAY_GPU_CALL(spraygpu_domain_bind(device_, id, scene_));
    }
    cache_block_ = cache_block;
    sinfo->rtc_scene = scene_;
    sinfo->cache_block = cache_block;
  }
  bool isResident(int id) const { return cache_.isLoaded(id); }
  std::size_t cacheHits() const { return cache_hits_; }
  std::size_t cacheMisses() const { return cache_misses_; }
  int cacheCapacity() const { return cache_.getCacheSize(); }

  // ---- per-ray forms (scene.h:157-195): a batch of one through the C ABI
  bool intersect(spraygpu_scene rtc_scene, int cache_block, const float org[3], const float dir[3],
                 RTCRayIntersection* isect) const {
    makeRadianceRay(org, dir, isect);
    const spraygpu_ray r = {org[0], org[1], org[2], isect->tnear, dir[0], dir[1], dir[2], INFINITY};
    spraygpu_key key;
    float uv[2];
    const int dom = block_domain_[(std::size_t)cache_block];
    SPRAY_GPU_CALL(spraygpu_intersect(rtc_scene, dom, &r, 1, &key, uv));
    if (key == SPRAYGPU_KEY_MISS) return false;
    const uint32_t tb = (uint32_t)(key >> 32);
    std::memcpy(&isect->tfar, &tb, 4);
    isect->u = uv[0]; isect->v = uv[1];
    isect->geomID = 0; isect->primID = (unsigned)(key & 0x7fffffu);
    surface_buf_.updateIntersection(cache_block, isect);
    return true;
  }
  bool intersect(const float org[3], const float dir[3], RTCRayIntersection* isect) const {
    return intersect(scene_, cache_block_, org, dir, isect);
  }
  bool occluded(spraygpu_scene rtc_scene, const float org[3], const float dir[3]) const {
    const spraygpu_ray r = {org[0], org[1], org[2], SPRAY_RAY_EPSILON, dir[0], dir[1], dir[2], INFINITY};
    uint8_t occ = 0;
    SPRAY_GPU_CALL(spraygpu_occluded(rtc_scene, &r, 1, &occ));
    return occ != 0;
  }
  bool occluded(const float org[3], const float dir[3]) const { return occluded(scene_, org, dir); }

  // ---- batched forms: n rays against ONE domain (host or device pointers)
  void intersect(spraygpu_scene rtc_scene, int domain_id, const spraygpu_ray* rays, std::size_t n, spraygpu_key* keys,
                 float* uv) const {
    SPRAY_GPU_CALL(spraygpu_intersect(rtc_scene, domain_id, rays, n, keys, uv));
  }
  void occluded(spraygpu_scene rtc_scene, const spraygpu_ray* rays, std::size_t n, uint8_t* occ) const {
    SPRAY_GPU_CALL(spraygpu_occluded(rtc_scene, rays, n, occ));
  }
  // ray -> sorted list of entered domains (Scene::intersectDomains -> WbvhEmbree::intersect, scene.h:197)
  void intersectDomains(const spraygpu_ray* rays, std::size_t n, int cap, int32_t* counts, int32_t* ids, float* tmins) const {
    SPRAY_GPU_CALL(spraygpu_domain_lists(device_, rays, n, cap, counts, ids, tmins));
  }
  void updateIntersection(RTCRayIntersection* isect) const { surface_buf_.updateIntersection(cache_block_, isect); }

  const Bsdf* getBsdf(int id) const { return &domains_[(std::size_t)id].bsdf; }
  std::size_t getNumDomains() const { return domains_.size(); }
  const std::vector<Domain>& getDomains() const { return domains_; }
  const std::vector<Light*>& getLights() const { return lights_; }
  std::size_t getNumLights() const { return lights_.size(); }

 private:
  // RTCRayUtil::makeRadianceRay (src/render/rays.h:345-363)
  static void makeRadianceRay(const float org[3], const float dir[3], RTCRayIntersection* r) {
    for (int k = 0; k < 3; ++k) { r->org[k] = org[k]; r->dir[k] = dir[k]; }
    r->tnear = SPRAY_RAY_EPSILON; r->tfar = INFINITY;
    r->geomID = RTC_INVALID_GEOMETRY_ID; r->primID = RTC_INVALID_GEOMETRY_ID;
    r->u = r->v = 0.f; r->color = 0; r->Ns[0] = r->Ns[1] = r->Ns[2] = 0.f;
  }

  // scene.inl:296-349
  void mergeDomainBounds(std::size_t* max_num_vertices, std::size_t* max_num_faces) {
    Aabb world;
    std::size_t nv = 0, nf = 0;
    for (Domain& d : domains_) {
      SPRAY_CHECK(d.world_aabb.isValid(), "domain " << d.id << ": `bound` missing or inverted");
      SPRAY_CHECK(d.num_vertices > 0, "domain " << d.id << ": `vertex` missing");
      SPRAY_CHECK(d.num_faces > 0, "domain " << d.id << ": `face` missing");
      if (d.num_vertices > nv) nv = d.num_vertices;
      if (d.num_faces > nf) nf = d.num_faces;
      world.merge(d.world_aabb);
    }
    *max_num_vertices = nv; *max_num_faces = nf;
    bound_ = world;
    SPRAY_CHECK(bound_.isValid(), "");
  }

  Aabb bound_;
  std::vector<Domain> domains_;
  std::vector<Light*> lights_;
  CacheT cache_;
  mutable SurfaceBufT surface_buf_;
  spraygpu_scene scene_ = nullptr;  // current domain's scene
  int cache_block_ = 0;             // current cache block
  std::vector<int> block_domain_;
  InsituPartition partition_;
  bool insitu_ = false;
  spraygpu_device device_ = nullptr;
  bool owns_device_ = false;
  std::size_t cache_hits_ = 0, cache_misses_ = 0;
};

}  // namespace spray
