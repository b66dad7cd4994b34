// gpu_tracer.h — GpuTracer<SceneT, ShaderTag>: the `TracerT` concept SprayRenderer consumes
// (reference src/render/spray_renderer.h:45-77; ooc::Tracer, src/ooc/ooc_tracer.h:50-63,
// ooc_tracer.inl:36-234; insitu::{Single,Multi}ThreadTracer, src/insitu/insitu_singlethread_tracer.inl:381-466):
//     typedef SceneType;  init(cfg, camera, scene*, image*);  trace();  traceInOmp();  type()
// One blocking tile = one call of the fused persistent kernel (spraygpu_trace_tile) instead of
// genSingleEyes -> isectDomsRads -> PContext::isectPrims.  The frame is accumulated in a device frame
// buffer; HdrImage::composite() is the NCCL reduce / row-band gather, and the image reaches the pinned host
// HdrImage once per frame.
//   --partition image  (spray_ooc):    every rank holds all domains, renders its vertical stripe
//   --partition insitu (spray_insitu): domains Morton-partitioned, spraygpu_trace_tile_insitu on full tiles
#pragma once
#include <cuda_runtime_api.h>

#include <vector>

#include "spray/check.h"
#include "spray/comm.h"
#include "spray/config.h"
#include "spray/render_types.h"
#include "spray/scene.h"
#include "spray_gpu.h"

namespace spray {

enum TracerType { TRACER_TYPE_SPRAY_OOC = 0, TRACER_TYPE_SPRAY_INSITU_1_THREAD, TRACER_TYPE_SPRAY_INSITU_N_THREADS };

struct ShaderPtTag { static constexpr int kind = SPRAYGPU_SHADER_PT; };  // ooc::ShaderPt / insitu::ShaderPt
struct ShaderAoTag { static constexpr int kind = SPRAYGPU_SHADER_AO; };  // ooc::ShaderAo / insitu::ShaderAo

template <typename SceneT, typename ShaderTag>
class GpuTracer {
 public:
  typedef SceneT SceneType;

  ~GpuTracer() {
    if (d_fb_) cudaFree(d_fb_);
    if (comm_) spraygpu_comm_destroy(comm_);
  }

  int type() const { return insitu_ ? TRACER_TYPE_SPRAY_INSITU_1_THREAD : TRACER_TYPE_SPRAY_OOC; }

  void init(const Config& cfg, const Camera& camera, SceneType* scene, HdrImage* image) {
    camera_ = &camera; scene_ = scene; image_ = image;
    rank_ = mpi::rank(); num_ranks_ = mpi::size();
    insitu_ = scene->insitu();
    SPRAY_CHECK(scene->getNumDomains() > 0, "");
    SPRAY_CHECK(cfg.pixel_samples > 0, "");
    SPRAY_CHECK(cfg.bounces > 0, "");
    SPRAY_CHECK(cfg.image_w > 0 && cfg.image_h > 0, "");
    dev_ = scene->device();

    // ShaderPt::init / ShaderAo::init (ooc_shader_pt.h:45-52): bounces, ao_samples, ks, shininess, lights
    shade_ = spraygpu_shade();
    shade_.shader = ShaderTag::kind;
    shade_.bounces = cfg.bounces;
    shade_.pixel_samples = cfg.pixel_samples;
    shade_.ao_samples = cfg.ao_samples;
    for (int k = 0; k < 3; ++k) shade_.ks[k] = cfg.ks[k];
    shade_.shininess = cfg.shininess;
    // in-situ tracers jitter multi-sample eye rays with (pixid, s), the ooc tracer with bufid (insitu_ray.h:161-165)
    shade_.eye_seed = insitu_ ? SPRAYGPU_EYE_SEED_INSITU : SPRAYGPU_EYE_SEED_OOC;
    const std::vector<Light*>& lights = scene->getLights();
    if (ShaderTag::kind == SPRAYGPU_SHADER_PT) SPRAY_CHECK(lights.size() > 0, "the PT shader needs a light (ooc_tracer.inl:68)");
    SPRAY_CHECK(lights.size() <= SPRAYGPU_MAX_LIGHTS, lights.size() << " lights");
    shade_.nlights = (int)lights.size();
    for (std::size_t i = 0; i < lights.size(); ++i) {
      shade_.lights[i].type = lights[i]->type;
      for (int k = 0; k < 3; ++k) { shade_.lights[i].position[k] = lights[i]->position[k]; shade_.lights[i].radiance[k] = lights[i]->radiance[k]; }
    }

    // image-partition ranks own a vertical stripe; in-situ ranks walk every tile together
    tile_list_.init(cfg.image_w, cfg.image_h, cfg.pixel_samples, insitu_ ? 1 : num_ranks_, insitu_ ? 0 : rank_,
                    cfg.maximum_num_screen_space_samples_per_rank);
    SPRAY_CHECK(insitu_ || !tile_list_.empty() || num_ranks_ > cfg.image_w, "");

    // --cache-size k < #domains: the out-of-core sweep (spraygpu_trace_tile_ooc) with Scene::load as its
    // residency callback, the reference's `scene->load(id)` inside PContext::isectPrims (ooc_pcontext.h:145-154)
    out_of_core_ = !insitu_ && scene->cacheCapacity() < (int)scene->getNumDomains();

    npixels_ = (std::size_t)cfg.image_w * cfg.image_h;
    SPRAY_CHECK(cudaSetDevice(cfg.cuda_device) == cudaSuccess, "cudaSetDevice");
    SPRAY_CHECK(cudaMalloc((void**)&d_fb_, npixels_ * 16) == cudaSuccess, "frame buffer allocation");
    if (num_ranks_ > 1) comm_ = mpi::createComm(dev_);
    image_->setCompositor([this]() { this->composite(); });
  }

  // one frame: SprayRenderer clears the image, then calls trace(), then image.composite() if clustered
  void trace() {
    const spraygpu_camera cam = camera_->toGpu();
    // The reference loads domains on demand inside PContext::isectPrims (ooc_pcontext.h:145-154); the fused
    // kernel wants them resident before the tile starts.  With an LRU cache that holds every domain this is
    // the first frame's one-time load (InfiniteCache scenes were warmed in Scene::init).
    if (!insitu_ && !out_of_core_)
      for (std::size_t id = 0; id < scene_->getNumDomains(); ++id)
        if (!scene_->isResident((int)id)) scene_->load((int)id);
    SPRAY_CHECK(cudaMemsetAsync(d_fb_, 0, npixels_ * 16, 0) == cudaSuccess, "");  // image_.clear() on the device
    if (!out_of_core_ && !(insitu_ && num_ranks_ > 1)) {
      // resident scene: the whole tile loop in one persistent launch
      std::vector<spraygpu_tile> gts;
      for (const Tile& t : tile_list_.tiles()) gts.push_back(spraygpu_tile{t.x, t.y, t.w, t.h});
      SPRAY_GPU_CALL(spraygpu_trace_frame(dev_, &cam, &shade_, gts.data(), (int)gts.size(), d_fb_, nullptr));
      if (num_ranks_ == 1) download();
      return;
    }
    for (const Tile& t : tile_list_.tiles()) {
      const spraygpu_tile gt = {t.x, t.y, t.w, t.h};
      if (out_of_core_)
        SPRAY_GPU_CALL(spraygpu_trace_tile_ooc(dev_, &cam, &shade_, gt, d_fb_, scene_->cacheCapacity(), &GpuTracer::loadDomain,
                                               scene_, nullptr));
      else if (insitu_ && num_ranks_ > 1) SPRAY_GPU_CALL(spraygpu_trace_tile_insitu(dev_, comm_, &cam, &shade_, gt, d_fb_, nullptr));
      else SPRAY_GPU_CALL(spraygpu_trace_tile(dev_, &cam, &shade_, gt, d_fb_, nullptr));
    }
    if (num_ranks_ == 1) download();
  }
  void traceInOmp() { trace(); }  // the reference's --dev-mode variant; there is no OpenMP team here

 private:
  // spraygpu_load_fn: Scene::load (LRU hit, or PLY parse + commit + bind on a miss)
  static int loadDomain(void* user, int domain_id) {
    static_cast<SceneType*>(user)->load(domain_id);
    return 0;
  }
  void composite() {
    if (num_ranks_ == 1) return;
    if (insitu_) {
      SPRAY_GPU_CALL(spraygpu_image_reduce(dev_, comm_, d_fb_, npixels_, 0));
    } else {
      // disjoint vertical stripes: a sum is exact (the other ranks contribute zeros)
      SPRAY_GPU_CALL(spraygpu_image_reduce(dev_, comm_, d_fb_, npixels_, 0));
    }
    if (rank_ == 0) download();
    else SPRAY_GPU_CALL(spraygpu_device_synchronize(dev_));
  }
  void download() {
    SPRAY_CHECK(cudaMemcpyAsync(image_->buf, d_fb_, npixels_ * 16, cudaMemcpyDeviceToHost, 0) == cudaSuccess, "");
    SPRAY_CHECK(cudaStreamSynchronize(0) == cudaSuccess, "frame failed on the device");
  }

  const Camera* camera_ = nullptr;
  SceneType* scene_ = nullptr;
  HdrImage* image_ = nullptr;
  spraygpu_device dev_ = nullptr;
  spraygpu_comm comm_ = nullptr;
  spraygpu_shade shade_;
  ImageScheduleTileList tile_list_;
  float* d_fb_ = nullptr;
  std::size_t npixels_ = 0;
  int rank_ = 0, num_ranks_ = 1;
  bool insitu_ = false;
  bool out_of_core_ = false;
};

}  // namespace spray
