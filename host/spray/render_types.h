// render_types.h — Tile / ImageScheduleTileList (reference src/render/tile.h, tile.cc:204-235,317-391),
// Camera (src/render/camera.h:128-209) and HdrImage (src/display/image.h:58-239) for the film-mode host
// driver.  Integer tiling logic and the float camera frame are restated with the reference's operation
// order; the frame buffer lives in pinned host memory so that the C ABI's D2H copy is a straight DMA.
#pragma once
#include <algorithm>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include "spray/check.h"
#include "spray/vecmath.h"
#include "spray_gpu.h"

namespace spray {

struct Tile {
  int x = 0, y = 0, w = 0, h = 0;
  int getArea() const { return w * h; }
};

inline Tile makeVerticalStripe(int num_ranks, int rank, const Tile& in) {
  Tile out;
  const int w = std::max(in.w / num_ranks, 1);
  out.x = in.x + rank * w;
  const int xend = in.x + in.w;
  if (out.x >= xend) { out.w = 0; out.h = 0; return out; }
  out.w = ((out.x + w > xend) || (rank == num_ranks - 1)) ? xend - out.x : w;
  out.y = in.y; out.h = in.h;
  return out;
}

inline Tile makeHorizontalStripe(int num_ranks, int rank, const Tile& in) {
  Tile out;
  const int h = std::max(in.h / num_ranks, 1);
  out.y = in.y + rank * h;
  const int yend = in.y + in.h;
  if (out.y >= yend) { out.w = 0; out.h = 0; return out; }
  out.h = ((out.y + h > yend) || (rank == num_ranks - 1)) ? yend - out.y : h;
  out.x = in.x; out.w = in.w;
  return out;
}

// this rank's vertical stripe cut into blocking tiles of <= maximum_num_samples_per_rank samples
class ImageScheduleTileList {
 public:
  void init(int64_t image_w, int64_t image_h, int64_t num_pixel_samples, int64_t num_ranks, int rank,
            int64_t maximum_num_samples_per_rank) {
    tiles_.clear();
    largest_ = 0;
    Tile image; image.w = (int)image_w; image.h = (int)image_h;
    const Tile vs = makeVerticalStripe((int)num_ranks, rank, image);
    if (vs.getArea() > 0) {
      const int64_t num_samples = (int64_t)vs.w * (int64_t)vs.h * num_pixel_samples;
      const int64_t est = (num_samples + maximum_num_samples_per_rank - 1) / maximum_num_samples_per_rank;
      SPRAY_CHECK(est < INT_MAX, "");
      const int tile_h = std::max(vs.h / (int)est, 1);  // the reference divides without the guard
      int max_area = -1;
      for (int y = 0; y < vs.h; y += tile_h) {
        Tile t; t.x = vs.x; t.y = y; t.w = vs.w; t.h = std::min(tile_h, vs.h - y);
        if (t.getArea() > max_area) { max_area = t.getArea(); largest_ = tiles_.size(); }
        tiles_.push_back(t);
      }
    }
    index_ = 0;
  }
  void reset() { index_ = 0; }
  const Tile& front() const { return tiles_[index_]; }
  void pop() { ++index_; }
  bool empty() const { return index_ == tiles_.size(); }
  std::size_t size() const { return tiles_.size(); }
  const Tile& getLargestBlockingTile() const { return tiles_[largest_]; }
  const std::vector<Tile>& tiles() const { return tiles_; }

 private:
  std::vector<Tile> tiles_;
  std::size_t index_ = 0, largest_ = 0;
};

class Camera {
 public:
  void init(vec3 campos, vec3 lookat, vec3 upvec, float vfov, int image_w, int image_h) {
    const float aspect = (float)image_w / (float)image_h;
    // `vfov * M_PI / 180.0f` is evaluated in double and narrowed (camera.h:133)
    const float theta = (float)((double)vfov * M_PI / (double)180.0f);
    const float half_h = (float)std::tan((double)(theta / 2.0f));
    const float half_w = aspect * half_h;
    vec3 w = campos - lookat;
    vec3 u = cross(upvec, w);
    vec3 v = cross(w, u);
    w = normalize(w); u = normalize(u); v = normalize(v);
    const vec3 center = campos - w;
    lowerleft_ = center - (half_w * u) - (half_h * v);
    wvec_ = (2.0f * half_w) * u;
    hvec_ = (2.0f * half_h) * v;
    pos_ = campos; lookat_ = lookat; up_ = upvec; vfov_ = vfov;
    image_w_ = image_w; image_h_ = image_h;
  }
  vec3 getPosition() const { return pos_; }
  vec3 getLookAt() const { return lookat_; }
  vec3 getUp() const { return up_; }
  // eye-ray direction of pixel-space point (x, y): camera.h:171-209
  vec3 getDirection(float x, float y) const {
    const float u = x / (float)image_w_, v = y / (float)image_h_;
    return normalize(lowerleft_ + (u * wvec_) + (v * hvec_) - pos_);
  }
  spraygpu_camera toGpu() const {
    spraygpu_camera c;
    for (int k = 0; k < 3; ++k) { c.pos[k] = pos_[k]; c.lowerleft[k] = lowerleft_[k]; c.wvec[k] = wvec_[k]; c.hvec[k] = hvec_[k]; }
    c.image_w = image_w_; c.image_h = image_h_;
    return c;
  }

 private:
  vec3 pos_, lookat_, up_, lowerleft_, wvec_, hvec_;
  float vfov_ = 90.f;
  int image_w_ = 0, image_h_ = 0;
};

void* allocPinned(std::size_t bytes);  // cudaHostAlloc, falls back to aligned malloc without a device
void freePinned(void* p);

struct HdrImage {
  HdrImage() {}
  ~HdrImage() { freePinned(buf); }
  HdrImage(const HdrImage&) = delete;
  HdrImage& operator=(const HdrImage&) = delete;

  void resize(int width, int height) {
    freePinned(buf);
    buf = (float*)allocPinned((std::size_t)width * height * 4 * sizeof(float));
    w = width; h = height;
    clear();
  }
  int bytes() const { return w * h * 4 * (int)sizeof(float); }
  void clear() { std::memset(buf, 0, (std::size_t)w * h * 4 * sizeof(float)); }
  // HdrImage::add(int, const float[3], double) (image.h:99-109): the product is formed in double
  void add(int pixid, const float rgb[3], double scale) {
    float* p = buf + 4 * (std::size_t)pixid;
    p[0] += (float)(scale * rgb[0]); p[1] += (float)(scale * rgb[1]); p[2] += (float)(scale * rgb[2]);
  }
  // HdrImage::composite (image.h:167-181) = MPI_Reduce(SUM) to rank 0 in the reference; here the tracer
  // installs the NCCL reduction of the device-resident frame buffer
  void setCompositor(std::function<void()> f) { compositor_ = std::move(f); }
  void composite() { if (compositor_) compositor_(); }
  // P3 PPM, 10 bits per channel, rows bottom-up (image.h:183-204)
  void writePpm(const char* filename) const { writePpm(w, h, filename, buf); }
  static void writePpm(int width, int height, const char* filename, const void* rgba) {
    FILE* f = std::fopen(filename, "w");
    SPRAY_CHECK(f != nullptr, filename);
    std::fprintf(f, "P3\n%d %d\n1023\n", width, height);
    const float* b = (const float*)rgba;
    std::string out;
    out.reserve((std::size_t)width * 16);
    char tmp[48];
    for (int y = height - 1; y > -1; --y) {
      out.clear();
      for (int x = 0; x < width; ++x) {
        const std::size_t i = ((std::size_t)y * width + x) << 2;
        const float r = std::fmin(std::fmax(b[i] * 1023.f, 0.0f), 1023.0f);
        const float g = std::fmin(std::fmax(b[i + 1] * 1023.f, 0.0f), 1023.0f);
        const float bl = std::fmin(std::fmax(b[i + 2] * 1023.f, 0.0f), 1023.0f);
        const int n = std::snprintf(tmp, sizeof(tmp), "%u %u %u\n", (unsigned)r, (unsigned)g, (unsigned)bl);
        out.append(tmp, (std::size_t)n);
      }
      std::fwrite(out.data(), 1, out.size(), f);
    }
    std::fclose(f);
  }

  int w = 0, h = 0;
  float* buf = nullptr;  // w*h float4, pinned

 private:
  std::function<void()> compositor_;
};

}  // namespace spray
