// spray_gpu — film-mode driver with the reference's command line (src/apps/spray_ooc.cc:30-112,
// src/apps/spray_insitu_singlethread.cc): same flags, same renderer / tracer / scene / cache type matrix,
// the Embree+OpenMP tracers replaced by the GPU tracer.
//
//   spray_gpu --mode film -w 2048 -h 2048 --frames 10 --cache-size -1 --partition image ...
//             --camera 90.17218 84.141418 82.480225 30 28.649426 30 --pixel-samples 1 --bounces 1 ...
//             --blinn 0.4 0.4 0.4 10 --ply-path examples/wavelets64 examples/wavelets64/wavelets64.spray
//
// Multi-GPU: one process per GPU (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_PORT in the environment, e.g.
// under `python -m torch.distributed.run --no-python`), --partition image or insitu.
#include <cstdio>

#include "spray/caches.h"
#include "spray/comm.h"
#include "spray/config.h"
#include "spray/gpu_tracer.h"
#include "spray/scene.h"
#include "spray/spray_renderer.h"

int main(int argc, char** argv) {
  spray::mpi::init();  // MPI_Init_thread + MPI_Comm_rank/size in the reference

  typedef spray::InfiniteCache InfCacheT;
  typedef spray::LruCache LruCacheT;
  typedef spray::Scene<InfCacheT> SceneInfT;
  typedef spray::Scene<LruCacheT> SceneLruT;

  typedef spray::SprayRenderer<spray::GpuTracer<SceneInfT, spray::ShaderAoTag>> RenderAoInfT;
  typedef spray::SprayRenderer<spray::GpuTracer<SceneLruT, spray::ShaderAoTag>> RenderAoLruT;
  typedef spray::SprayRenderer<spray::GpuTracer<SceneInfT, spray::ShaderPtTag>> RenderPtInfT;
  typedef spray::SprayRenderer<spray::GpuTracer<SceneLruT, spray::ShaderPtTag>> RenderPtLruT;

  spray::Config cfg;
  cfg.parse(argc, argv);

  if (cfg.partition == spray::Config::IMAGE) {  // spray_ooc.cc:78-104
    if (cfg.ao_mode) {
      if (cfg.cache_size < 0) { RenderAoInfT r; r.init(cfg); r.run(); }
      else { RenderAoLruT r; r.init(cfg); r.run(); }
    } else {
      if (cfg.cache_size < 0) { RenderPtInfT r; r.init(cfg); r.run(); }
      else { RenderPtLruT r; r.init(cfg); r.run(); }
    }
  } else if (cfg.partition == spray::Config::INSITU) {  // spray_insitu_*.cc: InfiniteCache only
    if (cfg.ao_mode) { RenderAoInfT r; r.init(cfg); r.run(); }
    else { RenderPtInfT r; r.init(cfg); r.run(); }
  } else {
    SPRAY_FATAL("unsupported partition " << cfg.partition);
  }
  return 0;
}
