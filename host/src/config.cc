// config.cc — see spray/config.h (reference src/render/config.cc:33-283).
#include "spray/config.h"

#include <getopt.h>

#include <cstdio>
#include <cstdlib>

#include "spray/check.h"

namespace spray {

Config::Config() {
  image_w = 400; image_h = 400;
  has_camera_config = false;
  camera_up = vec3(0.f, 1.f, 0.f);
  znear = 0.000001f; zfar = 1000000.0f; fov = 90.0f;
  nframes = -1;
  output_filename = "spray.ppm";
  light_samples = 1;
  bounces = 1;
  partition = IMAGE;
  num_partitions = 1;
  view_mode = VIEW_MODE_GLFW;
  cache_size = -1;
  ao_samples = 8;
  ao_mode = 0;
  pixel_samples = 4;
  maximum_num_screen_space_samples_per_rank = 1024 * 1024;
  nthreads = 1;
  shading = SPRAY_SHADING_LAMBERT;
  shininess = 0.f;
  ks = vec3(0.f, 0.f, 0.f);
  dev_mode = DEVMODE_NORMAL;
  const char* lr = std::getenv("LOCAL_RANK");
  cuda_device = lr ? std::atoi(lr) : 0;
}

void Config::printUsage(const char* argv0) {
  std::printf("Usage: %s [options] <modelfile>\n", argv0);
  std::printf("Options:\n  (): default setting\n");
  std::printf("  --output, -o <filename (spray.ppm)>\n     only effective in film mode\n");
  std::printf("  --local-disk, -l <path to local disk>\n");
  std::printf("  --ply-path <path to ply files>\n");
  std::printf("  --mode, -m <film | glfw | domain | partition>   (only film runs here: no graphics stack)\n");
  std::printf("  --num-partitions <number of partitions>, effective in partition view mode\n");
  std::printf("  --partition <image | hybrid | insitu>\n");
  std::printf("  --cache-size <max. number of domains>\n");
  std::printf("  --width, -w <image_width>\n  --height, -h <image_height>\n");
  std::printf("  --frames <number of frames (-1)>\n");
  std::printf("  --light-samples, <number of light samples (1)>\n");
  std::printf("  --bounces, <number of bounces (1)>\n");
  std::printf("  --camera-up <upx upy upz>\n  --camera <posx posy posz lookx looky lookz>\n");
  std::printf("  --fov <vertical field of view in degrees (90)>\n");
  std::printf("  --ao-samples <number of samples in AO (8)>\n  --ao-mode\n");
  std::printf("  --pixel-samples <number of pixel samples (4)>\n");
  std::printf("  --max-samples-per-rank <maximum number of screen-space samples per rank (1048576)>\n");
  std::printf("  --nthreads <number of threads (1)>   (accepted, unused: the GPU replaces the OpenMP team)\n");
  std::printf("  --shading <lambert | blinn>\n  --blinn ks_r ks_g ks_b shininess\n  --dev-mode\n");
  std::printf("  --cuda-device <n>   (this implementation; default LOCAL_RANK or 0)\n");
}

// util::parseTuple (src/utils/util.h:45-51): optarg is the first value, the rest follow in argv
static void parseTuple(char** argv, int argc, int len, float* out) {
  const int index = optind - 1;
  SPRAY_CHECK(index + len <= argc, "option needs " << len << " values");
  for (int i = 0; i < len; ++i) out[i] = (float)std::atof(argv[index + i]);
  optind = index + len;
}

void Config::parse(int argc, char** argv) {
  static struct option long_options[] = {
      {"output", required_argument, 0, 'o'},        {"local-disk", required_argument, 0, 'l'},
      {"mode", required_argument, 0, 'm'},          {"width", required_argument, 0, 'w'},
      {"height", required_argument, 0, 'h'},        {"frames", required_argument, 0, 300},
      {"nthreads", required_argument, 0, 302},      {"camera-up", required_argument, 0, 308},
      {"light-samples", required_argument, 0, 309}, {"bounces", required_argument, 0, 310},
      {"camera", required_argument, 0, 311},        {"cache-size", required_argument, 0, 313},
      {"fov", required_argument, 0, 314},           {"partition", required_argument, 0, 324},
      {"num-partitions", required_argument, 0, 325}, {"ao-samples", required_argument, 0, 400},
      {"ao-mode", no_argument, 0, 402},             {"pixel-samples", required_argument, 0, 403},
      {"shading", required_argument, 0, 404},       {"blinn", required_argument, 0, 405},
      {"max-samples-per-rank", required_argument, 0, 406}, {"ply-path", required_argument, 0, 408},
      {"dev-mode", no_argument, 0, 1000},           {"cuda-device", required_argument, 0, 2000},
      {0, 0, 0, 0}};
  optind = 1;  // allow repeated parses (tests)
  int option_index = 0, c;
  camera_up = vec3(0.f, 1.f, 0.f);
  while ((c = getopt_long(argc, argv, "o:l:t:m:r:w:h:x", long_options, &option_index)) != -1) {
    switch (c) {
      case 'o': output_filename = optarg; break;
      case 'l': local_disk_path = optarg; break;
      case 'm': {
        const std::string mode = optarg;
        if (mode == "domain") view_mode = VIEW_MODE_DOMAIN;
        else if (mode == "partition") view_mode = VIEW_MODE_PARTITION;
        else if (mode == "film") view_mode = VIEW_MODE_FILM;
        else if (mode == "glfw") view_mode = VIEW_MODE_GLFW;
        else SPRAY_FATAL("undefined view mode" << optarg);
      } break;
      case 'w': image_w = std::atoi(optarg); break;
      case 'h': image_h = std::atoi(optarg); break;
      case 300: nframes = std::atoi(optarg); break;
      case 302: nthreads = std::atoi(optarg); break;
      case 308: { float d[3]; parseTuple(argv, argc, 3, d); camera_up = vec3(d[0], d[1], d[2]); } break;
      case 309: light_samples = std::atoi(optarg); break;
      case 310: bounces = std::atoi(optarg); break;
      case 311: {
        has_camera_config = true;
        float d[6]; parseTuple(argv, argc, 6, d);
        camera_pos = vec3(d[0], d[1], d[2]); camera_lookat = vec3(d[3], d[4], d[5]);
      } break;
      case 313: cache_size = std::atoi(optarg); break;
      case 314: fov = (float)std::atof(optarg); break;
      case 324: {
        const std::string p = optarg;
        if (p == "image") partition = IMAGE;
        else if (p == "hybrid") partition = HYBRID;
        else if (p == "insitu") partition = INSITU;
        else SPRAY_FATAL("unsupported partitioning type: " << p);
      } break;
      case 325: num_partitions = std::atoi(optarg); break;
      case 400: ao_samples = std::atoi(optarg); break;
      case 402: ao_mode = 1; break;
      case 403: pixel_samples = std::atoi(optarg); break;
      case 404: shading = std::string(optarg) == "blinn" ? SPRAY_SHADING_BLINN : SPRAY_SHADING_LAMBERT; break;
      case 405: { float d[4]; parseTuple(argv, argc, 4, d); ks = vec3(d[0], d[1], d[2]); shininess = d[3]; } break;
      case 406: maximum_num_screen_space_samples_per_rank = std::atoi(optarg); break;
      case 408: ply_path = optarg; break;
      case 1000: dev_mode = DEVMODE_DEV; break;
      case 2000: cuda_device = std::atoi(optarg); break;
      default:
        printUsage(argv[0]);
        SPRAY_FATAL("undefined arg " << c);
    }
  }
  SPRAY_CHECK(argc != optind, "input file not found");
  model_descriptor_filename = argv[optind];
  const std::string ext = model_descriptor_filename.substr(model_descriptor_filename.find_last_of(".") + 1);
  SPRAY_CHECK(ext == "spray", "scene description must be a .spray file, got " << model_descriptor_filename);
}

}  // namespace spray
