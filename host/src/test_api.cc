// test_api.cc — extern "C" windows into the host-side logic for tests/test_host.py (ctypes).  Everything
// here runs without a GPU: PLY / .spray parsing, caches, tiles, camera frame, PPM writer, config parsing.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "spray/caches.h"
#include "spray/check.h"
#include "spray/config.h"
#include "spray/ply_loader.h"
#include "spray/render_types.h"
#include "spray/scene_loader.h"

using namespace spray;

#define API extern "C" __attribute__((visibility("default")))

static std::string g_err;
template <typename F>
static int guarded(F f) {
  setThrowOnFatal(true);
  try { f(); return 0; }
  catch (const FatalError& e) { g_err = e.what(); return -1; }
  catch (const std::exception& e) { g_err = e.what(); return -2; }
}
API const char* host_last_error() { return g_err.c_str(); }

API int host_ply_header(const char* path, uint64_t* nv, uint64_t* nf, int* has_color) {
  return guarded([&] {
    PlyLoader::Header h;
    PlyLoader::quickHeaderRead(path, &h);
    *nv = h.num_vertices; *nf = h.num_faces; *has_color = h.has_color ? 1 : 0;
  });
}

API int host_ply_load(const char* path, float* verts, uint64_t vcap_floats, uint32_t* faces, uint64_t fcap_idx,
                      uint32_t* colors, uint64_t ccap, uint64_t* nv, uint64_t* nf) {
  return guarded([&] {
    PlyLoader l;
    PlyLoader::Data d;
    d.vertices_capacity = vcap_floats; d.faces_capacity = fcap_idx; d.colors_capacity = ccap;
    d.vertices = verts; d.faces = faces; d.colors = colors;
    d.num_vertices = d.num_faces = 0;
    l.load(path, &d);
    *nv = d.num_vertices; *nf = d.num_faces;
  });
}

// scene file -> flat arrays.  per domain: object aabb 6, world aabb 6, transform 16 (column-major), mtl type,
// mtl params 3, num_vertices, num_faces  (34 doubles each); lights: type, pos 3, rad 3 (7 doubles each)
API int host_scene_load(const char* path, const char* ply_path, int max_domains, int max_lights, double* dom_out,
                        char* filenames /* max_domains x 256 */, double* light_out, int* ndomains, int* nlights) {
  return guarded([&] {
    std::vector<Domain> doms;
    std::vector<Light*> lights;
    SceneLoader sl;
    sl.load(path, ply_path ? ply_path : "", &doms, &lights);
    *ndomains = (int)doms.size(); *nlights = (int)lights.size();
    for (int i = 0; i < (int)doms.size() && i < max_domains; ++i) {
      const Domain& d = doms[(std::size_t)i];
      double* o = dom_out + 34 * i;
      for (int k = 0; k < 3; ++k) {
        o[k] = d.object_aabb.bounds[0][k]; o[3 + k] = d.object_aabb.bounds[1][k];
        o[6 + k] = d.world_aabb.bounds[0][k]; o[9 + k] = d.world_aabb.bounds[1][k];
      }
      for (int k = 0; k < 16; ++k) o[12 + k] = d.transform.data()[k];
      o[28] = d.bsdf.type;
      for (int k = 0; k < 3; ++k) o[29 + k] = d.bsdf.params[k];
      o[32] = (double)d.num_vertices; o[33] = (double)d.num_faces;
      std::strncpy(filenames + 256 * i, d.filename.c_str(), 255);
      filenames[256 * i + 255] = 0;
    }
    for (int i = 0; i < (int)lights.size(); ++i) {
      if (i < max_lights) {
        double* o = light_out + 7 * i;
        o[0] = lights[(std::size_t)i]->type;
        for (int k = 0; k < 3; ++k) { o[1 + k] = lights[(std::size_t)i]->position[k]; o[4 + k] = lights[(std::size_t)i]->radiance[k]; }
      }
      delete lights[(std::size_t)i];
    }
  });
}

// replay a load sequence; out[i] = block | (hit << 30)
API int host_cache_replay(int lru, int num_domains, int cache_size, const int* seq, int n, int* out, int* capacity) {
  return guarded([&] {
    if (lru) {
      LruCache c; c.init(num_domains, cache_size);
      *capacity = c.getCacheSize();
      for (int i = 0; i < n; ++i) { int b; const bool hit = c.load(seq[i], &b); out[i] = b | (hit ? 1 << 30 : 0); }
    } else {
      InfiniteCache c; c.init(num_domains, cache_size);
      *capacity = c.getCacheSize();
      for (int i = 0; i < n; ++i) { int b; const bool hit = c.load(seq[i], &b); out[i] = b | (hit ? 1 << 30 : 0); }
    }
  });
}

API int host_tiles(int w, int h, int spp, int nranks, int rank, int64_t max_samples, int max_tiles, int* xywh, int* ntiles,
                   int* largest_xywh) {
  return guarded([&] {
    ImageScheduleTileList tl;
    tl.init(w, h, spp, nranks, rank, max_samples);
    *ntiles = (int)tl.size();
    int i = 0;
    for (const Tile& t : tl.tiles()) {
      if (i < max_tiles) { xywh[4 * i] = t.x; xywh[4 * i + 1] = t.y; xywh[4 * i + 2] = t.w; xywh[4 * i + 3] = t.h; }
      ++i;
    }
    if (tl.size()) {
      const Tile& t = tl.getLargestBlockingTile();
      largest_xywh[0] = t.x; largest_xywh[1] = t.y; largest_xywh[2] = t.w; largest_xywh[3] = t.h;
    }
  });
}

API int host_stripes(int horizontal, int nranks, int rank, const int* in_xywh, int* out_xywh) {
  return guarded([&] {
    Tile in; in.x = in_xywh[0]; in.y = in_xywh[1]; in.w = in_xywh[2]; in.h = in_xywh[3];
    const Tile o = horizontal ? makeHorizontalStripe(nranks, rank, in) : makeVerticalStripe(nranks, rank, in);
    out_xywh[0] = o.x; out_xywh[1] = o.y; out_xywh[2] = o.w; out_xywh[3] = o.h;
  });
}

API int host_camera(const float* pos, const float* lookat, const float* up, float vfov, int w, int h, spraygpu_camera* out) {
  return guarded([&] {
    Camera c;
    c.init(vec3(pos[0], pos[1], pos[2]), vec3(lookat[0], lookat[1], lookat[2]), vec3(up[0], up[1], up[2]), vfov, w, h);
    *out = c.toGpu();
  });
}

API int host_write_ppm(const char* path, int w, int h, const float* rgba) {
  return guarded([&] { HdrImage::writePpm(w, h, path, rgba); });
}

// parses a command line; fills the numeric fields in a flat array (see tests/test_host.py for the order)
API int host_config_parse(int argc, char** argv, double* out, char* model, char* ply_path, char* output) {
  return guarded([&] {
    Config c;
    c.parse(argc, argv);
    int k = 0;
    out[k++] = c.image_w; out[k++] = c.image_h; out[k++] = c.has_camera_config;
    for (int a = 0; a < 3; ++a) out[k++] = c.camera_pos[a];
    for (int a = 0; a < 3; ++a) out[k++] = c.camera_lookat[a];
    for (int a = 0; a < 3; ++a) out[k++] = c.camera_up[a];
    out[k++] = c.fov; out[k++] = c.nframes; out[k++] = c.light_samples; out[k++] = c.bounces; out[k++] = c.partition;
    out[k++] = c.num_partitions; out[k++] = c.view_mode; out[k++] = c.cache_size; out[k++] = c.ao_samples; out[k++] = c.ao_mode;
    out[k++] = c.pixel_samples; out[k++] = c.maximum_num_screen_space_samples_per_rank; out[k++] = c.nthreads;
    out[k++] = c.shading; out[k++] = c.shininess;
    for (int a = 0; a < 3; ++a) out[k++] = c.ks[a];
    out[k++] = c.dev_mode;
    std::strncpy(model, c.model_descriptor_filename.c_str(), 255);
    std::strncpy(ply_path, c.ply_path.c_str(), 255);
    std::strncpy(output, c.output_filename.c_str(), 255);
  });
}
