// config.h — command-line configuration with the reference's flags, defaults and parsing rules
// (src/render/config.h:33-95, config.cc:33-283): getopt_long over the same option table, multi-value
// options (--camera, --camera-up, --blinn) consume the following argv words (util::parseTuple).
#pragma once
#include <string>

#include "spray/vecmath.h"

namespace spray {

enum ViewMode { VIEW_MODE_GLFW = 0, VIEW_MODE_FILM, VIEW_MODE_DOMAIN, VIEW_MODE_PARTITION, VIEW_MODE_TERMINATE };

class Config {
 public:
  Config();
  void parse(int argc, char** argv);
  static void printUsage(const char* argv0);

  int image_w, image_h;
  std::string model_descriptor_filename;
  std::string ply_path;

  bool has_camera_config;
  vec3 camera_pos, camera_lookat, camera_up;
  float znear, zfar, fov;

  int nframes;
  std::string output_filename;
  int light_samples;
  int bounces;

  enum Partition { IMAGE, HYBRID, INSITU };
  int partition;
  int num_partitions;

  int view_mode;
  int cache_size;

  int ao_samples;
  int ao_mode;
  int pixel_samples;
  int maximum_num_screen_space_samples_per_rank;

  std::string local_disk_path;
  int nthreads;

  enum Shading { SPRAY_SHADING_LAMBERT, SPRAY_SHADING_BLINN };
  int shading;
  float shininess;
  vec3 ks;  // uninitialised in the reference without --blinn; zero here

  enum DevMode { DEVMODE_NORMAL, DEVMODE_DEV };
  int dev_mode;

  // additions of this implementation (not reference flags)
  int cuda_device;   // --cuda-device <n> (default: LOCAL_RANK or 0)
};

}  // namespace spray
