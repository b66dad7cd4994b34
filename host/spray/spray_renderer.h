// spray_renderer.h — SprayRenderer<TracerT>: film-mode front end (reference
// src/render/spray_renderer.h:45-77, spray_renderer.inl:40-135,358-420).  init(cfg) builds the scene, the
// frame buffer, the camera and the tracer in the reference's order; run() renders cfg.nframes frames and
// rank 0 writes the PPM.  The GLFW / OpenGL modes (glfw, domain, partition views) are unavailable on this
// box (no graphics stack) and CHECK-fail with that message.
#pragma once
#include <chrono>
#include <cstdio>

#include "spray/check.h"
#include "spray/comm.h"
#include "spray/config.h"
#include "spray/render_types.h"
#include "spray/scene.h"

namespace spray {

template <class TracerT>
class SprayRenderer {
  typedef typename TracerT::SceneType SceneType;

 public:
  SprayRenderer() : cfg_(nullptr) {}

  void init(const Config& cfg) {
    cfg_ = &cfg;
    SPRAY_CHECK(cfg.view_mode == VIEW_MODE_FILM,
                "only --mode film runs here: the glfw/domain/partition views need OpenGL (unavailable on this box)");
    const bool insitu_mode = (cfg.partition == Config::INSITU);
    scene_.init(cfg.model_descriptor_filename, cfg.ply_path, cfg.local_disk_path, cfg.cache_size, cfg.view_mode,
                insitu_mode, cfg.num_partitions);
    scene_.buildWbvh();
    image_.resize(cfg.image_w, cfg.image_h);
    initCamera(cfg);
    tracer_.init(cfg, camera_, &scene_, &image_);
  }

  void run() {
    SPRAY_CHECK(cfg_ != nullptr, "failed to run. not configured.");
    renderFilm();  // --dev-mode (renderFilmInOmp) differs only in where the OpenMP team is opened
  }

  const HdrImage& image() const { return image_; }
  SceneType& scene() { return scene_; }
  double msPerFrame() const { return ms_per_frame_; }

 private:
  // spray_renderer.inl:87-105
  void initCamera(const Config& cfg) {
    vec3 campos, lookat, upvec;
    if (cfg.has_camera_config) {
      campos = cfg.camera_pos; lookat = cfg.camera_lookat; upvec = cfg.camera_up;
    } else {
      const Aabb aabb = scene_.getBound();
      const vec3 extent = aabb.getExtent();
      lookat = aabb.getCenter();
      campos = lookat + (length(extent) * 0.5f) * vec3(0.0f, 0.0f, 1.0f);
      upvec = vec3(0.0f, 1.0f, 0.0f);
    }
    camera_.init(campos, lookat, upvec, cfg.fov, cfg.image_w, cfg.image_h);
  }

  // spray_renderer.inl:358-384
  void renderFilm() {
    const bool cluster = mpi::size() > 1;
    const int64_t nframes = cfg_->nframes < 0 ? 1 : cfg_->nframes;  // -1 = "until closed" in glfw mode
    const auto t0 = std::chrono::steady_clock::now();
    for (int64_t i = 0; i < nframes; ++i) {
      image_.clear();
      if (cfg_->dev_mode == Config::DEVMODE_DEV) tracer_.traceInOmp();
      else tracer_.trace();
      if (cluster) image_.composite();
    }
    const auto t1 = std::chrono::steady_clock::now();
    ms_per_frame_ = std::chrono::duration<double, std::milli>(t1 - t0).count() / (double)(nframes > 0 ? nframes : 1);
    if (mpi::rank() == 0) {
      image_.writePpm(cfg_->output_filename.c_str());
      // the reference prints its timers through utils/profiler.cc:61-113 when built with SPRAY_TIMING
      std::printf("[STATS] total : %.3f ms/frame (%lld frames), frame rate %.2f fps\n", ms_per_frame_, (long long)nframes,
                  ms_per_frame_ > 0 ? 1000.0 / ms_per_frame_ : 0.0);
    }
  }

  const Config* cfg_;
  SceneType scene_;
  Camera camera_;
  TracerT tracer_;
  HdrImage image_;
  double ms_per_frame_ = 0.0;
};

}  // namespace spray
