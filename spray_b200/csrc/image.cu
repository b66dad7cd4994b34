// image.cu — the frame image of the GPU path: spraygpu_image.
//
// Reference: HdrImage (src/display/image.h:58-239) — `buf` (vec4 per pixel, host), clear(), add(), composite() =
// MPI_Reduce(SUM) of every rank's WHOLE image into rank 0 (image.h:167-181), writePpm().  SprayRenderer::renderFilm
// (src/render/spray_renderer.inl:355-372) per frame: image.clear(); tracer.trace(); image.composite().
//
// Here a frame is traced into a DEVICE render target (atomics land in L2).  What the reference does with a
// whole-image reduce is a FLUSH of the rows this rank rendered, done by one kernel of this library with plain 16-byte
// stores on peer-mapped / host-mapped pointers:
//     to_root : into the root rank's composite image over NVLink (CUDA IPC mapping; image partition: rows are owned
//               by exactly one rank, so a store replaces the reduce's sum)
//     to_host : into the pinned host image over PCIe (the application's HdrImage::buf; with a POSIX shared-memory
//               name all ranks of a box write ONE host image, each over its own PCIe link)
// and the same kernel zeroes the render target (next frame's image.clear()).  Only pixels that are non-zero now, or
// were non-zero when they were last flushed to that destination, are written ("sparse delta": a bit per pixel and
// destination remembers what the destination holds) — C2 touches 27 % of its pixels, so 17 of 64 MB cross the wire;
// the destination is bit-identical to a full copy as long as nobody else writes it.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <string>
#include <vector>

#include "internal.cuh"

namespace spraygpu {

struct Comm;
int comm_shared_alloc(Comm* c, size_t bytes, void** local, std::vector<void*>* mapped, int only_rank);
void comm_shared_free(Comm* c, void* local, std::vector<void*>* mapped);
int comm_host_barrier(Comm* c);
int comm_rank(const Comm* c);
int comm_size(const Comm* c);
Device* comm_device(const Comm* c);

struct Image {
  Device* dev = nullptr;
  Comm* comm = nullptr;
  int w = 0, h = 0, rank = 0, nranks = 1, root = 0;
  float4* d_target = nullptr;          // this rank's CURRENT render target (= d_targets[cur])
  // spraygpu_image_flush_async: two targets — frame k is flushed (and cleared) on `flush_stream` while frame k + 1 is
  // traced into the other one on the device's stream
  float4* d_targets[2] = {nullptr, nullptr};
  int cur = 0;
  cudaStream_t flush_stream = nullptr;
  cudaEvent_t ev_trace = nullptr, ev_flush[2] = {nullptr, nullptr};
  bool flushed_once[2] = {false, false};
  float4* d_root_local = nullptr;      // root: the composite image (peer-mapped into every rank); others: nullptr
  std::vector<void*> root_mapped;      // [nranks] mappings of every rank's shared buffer; [root] is the composite
  float4* h_image = nullptr;           // pinned host image (private cudaHostAlloc, or shared memory + cudaHostRegister)
  float4* h_image_dev = nullptr;       // the address kernels use for it (cudaHostGetDevicePointer)
  std::string shm_name;
  bool shm_owner = false;
  size_t bytes = 0;
  // rows this rank renders: blocks [row_begin[b], row_begin[b+1]) with owner[b] == rank
  std::vector<int32_t> my_begin, my_rows;
  int2* d_blocks = nullptr;            // (first row, row count) of my blocks
  int nmy = 0, max_rows = 0;
  uint32_t* d_prev_root = nullptr;     // 1 bit / pixel: the root composite holds a non-zero value there
  uint32_t* d_prev_host = nullptr;     // 1 bit / pixel: the host image holds a non-zero value there
  unsigned long long* d_flushed = nullptr;   // [0] pixels stored to root, [1] pixels stored to host (cumulative)
};

// One warp handles 32 consecutive pixels of a row block; the two "what the destination holds" bitmaps are read and
// written one word per warp.
__global__ void __launch_bounds__(256)
    image_flush_kernel(float4* __restrict__ target, float4* __restrict__ root, float4* __restrict__ host, uint32_t* __restrict__ prev_root,
                       uint32_t* __restrict__ prev_host, const int2* __restrict__ blocks, int image_w, int clear,
                       unsigned long long* __restrict__ flushed) {
  const int2 blk = blocks[blockIdx.y];
  const size_t npix = (size_t)blk.y * image_w;
  const size_t base = (size_t)blk.x * image_w;            // first pixel of the block: a multiple of 32 (image_w % 32 == 0)
  const int lane = threadIdx.x & 31;
  for (size_t p0 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) & ~(size_t)31; p0 < npix; p0 += (size_t)gridDim.x * blockDim.x) {
    const size_t p = p0 + lane;
    const bool in = p < npix;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in) v = target[base + p];
    const bool nz = in && (v.x != 0.0f || v.y != 0.0f || v.z != 0.0f || v.w != 0.0f);
    const uint32_t nzm = __ballot_sync(0xffffffffu, nz);
    const size_t word = (base + p0) >> 5;                 // one bitmap word per 32-pixel group of the image
    if (root) {
      uint32_t pm = 0u;
      if (lane == 0) pm = prev_root[word];
      pm = __shfl_sync(0xffffffffu, pm, 0);
      if (in && ((nzm | pm) >> lane & 1u)) root[base + p] = v;
      if (lane == 0) {
        if (pm != nzm) prev_root[word] = nzm;
        if (flushed && (nzm | pm)) atomicAdd(flushed + 0, (unsigned long long)__popc(nzm | pm));
      }
    }
    if (host) {
      uint32_t pm = 0u;
      if (lane == 0) pm = prev_host[word];
      pm = __shfl_sync(0xffffffffu, pm, 0);
      if (in && ((nzm | pm) >> lane & 1u)) host[base + p] = v;
      if (lane == 0) {
        if (pm != nzm) prev_host[word] = nzm;
        if (flushed && (nzm | pm)) atomicAdd(flushed + 1, (unsigned long long)__popc(nzm | pm));
      }
    }
    if (clear && nz) target[base + p] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

}  // namespace spraygpu

using namespace spraygpu;

extern "C" {

int spraygpu_image_create(spraygpu_device dev_h, spraygpu_comm comm_h, int image_w, int image_h, const char* shm_name,
                          spraygpu_image* out) {
  SG_CHECK(dev_h && out && image_w > 0 && image_h > 0, SPRAYGPU_ERR_ARG, "image_create: bad argument");
  SG_CHECK(image_w % 32 == 0, SPRAYGPU_ERR_ARG, "image_create: image width %d must be a multiple of 32 (one bitmap word per 32 pixels of a row)", image_w);
  *out = nullptr;
  Device* dev = (Device*)dev_h;
  Comm* comm = (Comm*)comm_h;
  SG_CHECK(!comm || comm_device(comm) == dev, SPRAYGPU_ERR_ARG, "image_create: communicator belongs to another device");
  SG_CUDA(cudaSetDevice(dev->cuda_device));
  Image* im = new Image();
  im->dev = dev; im->comm = comm; im->w = image_w; im->h = image_h;
  im->rank = comm ? comm_rank(comm) : 0; im->nranks = comm ? comm_size(comm) : 1;
  im->bytes = (size_t)image_w * image_h * sizeof(float4);
  const size_t words = ((size_t)image_w * image_h + 31) / 32;
  int rc = SPRAYGPU_OK;
  do {
    if (cudaMalloc((void**)&im->d_target, im->bytes) != cudaSuccess || cudaMalloc((void**)&im->d_prev_root, words * 4) != cudaSuccess ||
        cudaMalloc((void**)&im->d_prev_host, words * 4) != cudaSuccess || cudaMalloc((void**)&im->d_flushed, 16) != cudaSuccess) {
      set_error("image_create: device allocation failed: %s", cudaGetErrorString(cudaGetLastError())); rc = SPRAYGPU_ERR_CUDA; break;
    }
    im->d_targets[0] = im->d_target;
    cudaMemsetAsync(im->d_target, 0, im->bytes, dev->stream);
    cudaMemsetAsync(im->d_prev_root, 0, words * 4, dev->stream);
    cudaMemsetAsync(im->d_prev_host, 0, words * 4, dev->stream);
    cudaMemsetAsync(im->d_flushed, 0, 16, dev->stream);
    if (comm && im->nranks > 1) {
      // the composite lives on the root; the other ranks allocate a token buffer so that the exchange stays symmetric
      void* local = nullptr;
      rc = comm_shared_alloc(comm, im->rank == im->root ? im->bytes : 256, &local, &im->root_mapped, im->root);
      if (rc) break;
      im->d_root_local = (float4*)local;
    }
    // host image
    if (shm_name && shm_name[0]) {
      im->shm_name = shm_name[0] == '/' ? shm_name : std::string("/") + shm_name;
      int fd = -1;
      if (im->rank == im->root) {
        shm_unlink(im->shm_name.c_str());
        fd = shm_open(im->shm_name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd >= 0 && ftruncate(fd, (off_t)im->bytes) != 0) { close(fd); fd = -1; }
        im->shm_owner = fd >= 0;
      }
      if (comm) { rc = comm_host_barrier(comm); if (rc) break; }       // the segment exists before anyone opens it
      if (im->rank != im->root) fd = shm_open(im->shm_name.c_str(), O_RDWR, 0600);
      if (fd < 0) { set_error("image_create: shm_open(%s) failed", im->shm_name.c_str()); rc = SPRAYGPU_ERR_STATE; break; }
      void* p = mmap(nullptr, im->bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
      close(fd);
      if (p == MAP_FAILED) { set_error("image_create: mmap of %s failed", im->shm_name.c_str()); rc = SPRAYGPU_ERR_STATE; break; }
      if (im->rank == im->root) memset(p, 0, im->bytes);
      im->h_image = (float4*)p;
      if (cudaHostRegister(p, im->bytes, cudaHostRegisterPortable | cudaHostRegisterMapped) != cudaSuccess) {
        set_error("image_create: cudaHostRegister failed: %s", cudaGetErrorString(cudaGetLastError())); rc = SPRAYGPU_ERR_CUDA; break;
      }
      if (comm) { rc = comm_host_barrier(comm); if (rc) break; }       // zero-filled before anyone flushes into it
    } else {
      if (cudaHostAlloc((void**)&im->h_image, im->bytes, cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) {
        set_error("image_create: cudaHostAlloc failed: %s", cudaGetErrorString(cudaGetLastError())); rc = SPRAYGPU_ERR_CUDA; break;
      }
      memset(im->h_image, 0, im->bytes);
    }
    if (cudaHostGetDevicePointer((void**)&im->h_image_dev, im->h_image, 0) != cudaSuccess) {
      set_error("image_create: cudaHostGetDevicePointer failed: %s", cudaGetErrorString(cudaGetLastError())); rc = SPRAYGPU_ERR_CUDA; break;
    }
    if (cudaStreamSynchronize(dev->stream) != cudaSuccess) { set_error("image_create: %s", cudaGetErrorString(cudaGetLastError())); rc = SPRAYGPU_ERR_CUDA; break; }
  } while (0);
  if (rc) { spraygpu_image_destroy((spraygpu_image)im); return rc; }
  *out = (spraygpu_image)im;
  // default: one block = the whole image, rendered by this rank
  const int32_t rb[2] = {0, image_h};
  return spraygpu_image_set_rows(*out, 1, rb, nullptr);
}

int spraygpu_image_destroy(spraygpu_image img) {
  if (!img) return SPRAYGPU_OK;
  Image* im = (Image*)img;
  cudaSetDevice(im->dev->cuda_device);
  cudaStreamSynchronize(im->dev->stream);
  if (im->comm && !im->root_mapped.empty()) {
    comm_host_barrier(im->comm);       // nobody is still storing through a mapping that is about to go away
    comm_shared_free(im->comm, im->d_root_local, &im->root_mapped);
  }
  if (im->flush_stream) cudaStreamSynchronize(im->flush_stream);
  if (im->h_image) {
    if (!im->shm_name.empty()) {
      cudaHostUnregister(im->h_image);
      munmap(im->h_image, im->bytes);
      if (im->shm_owner) shm_unlink(im->shm_name.c_str());
    } else cudaFreeHost(im->h_image);
  }
  if (im->flush_stream) { cudaStreamSynchronize(im->flush_stream); cudaStreamDestroy(im->flush_stream); }
  if (im->ev_trace) cudaEventDestroy(im->ev_trace);
  for (int k = 0; k < 2; ++k) {
    if (im->ev_flush[k]) cudaEventDestroy(im->ev_flush[k]);
    if (im->d_targets[k]) cudaFree(im->d_targets[k]);
  }
  if (im->d_prev_root) cudaFree(im->d_prev_root);
  if (im->d_prev_host) cudaFree(im->d_prev_host);
  if (im->d_flushed) cudaFree(im->d_flushed);
  if (im->d_blocks) cudaFree(im->d_blocks);
  delete im;
  return SPRAYGPU_OK;
}

float* spraygpu_image_device(spraygpu_image img) { return img ? (float*)((Image*)img)->d_target : nullptr; }
float* spraygpu_image_host(spraygpu_image img) { return img ? (float*)((Image*)img)->h_image : nullptr; }
float* spraygpu_image_root(spraygpu_image img) {
  if (!img) return nullptr;
  Image* im = (Image*)img;
  if (im->nranks == 1) return (float*)im->d_target;
  return (float*)im->root_mapped[(size_t)im->root];
}

int spraygpu_image_set_rows(spraygpu_image img, int nblocks, const int32_t* row_begin, const int32_t* owner) {
  SG_CHECK(img && nblocks >= 0 && (row_begin || nblocks == 0), SPRAYGPU_ERR_ARG, "image_set_rows: bad argument");
  Image* im = (Image*)img;
  std::vector<int2> mine;
  int max_rows = 0;
  for (int b = 0; b < nblocks; ++b) {
    SG_CHECK(row_begin[b] >= 0 && row_begin[b] <= row_begin[b + 1] && row_begin[b + 1] <= im->h, SPRAYGPU_ERR_ARG,
             "image_set_rows: block %d = rows [%d, %d) of %d", b, row_begin[b], row_begin[b + 1], im->h);
    if (owner) SG_CHECK(owner[b] >= -1 && owner[b] < im->nranks, SPRAYGPU_ERR_ARG, "image_set_rows: block %d owner %d", b, owner[b]);
    if (owner && owner[b] != im->rank) continue;
    const int rows = row_begin[b + 1] - row_begin[b];
    if (rows <= 0) continue;
    mine.push_back(make_int2(row_begin[b], rows));
    max_rows = std::max(max_rows, rows);
  }
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  SG_CUDA(cudaStreamSynchronize(im->dev->stream));
  if (im->d_blocks) { cudaFree(im->d_blocks); im->d_blocks = nullptr; }
  im->nmy = (int)mine.size(); im->max_rows = max_rows;
  if (im->nmy) {
    SG_CUDA(cudaMalloc((void**)&im->d_blocks, sizeof(int2) * mine.size()));
    SG_CUDA(cudaMemcpy(im->d_blocks, mine.data(), sizeof(int2) * mine.size(), cudaMemcpyHostToDevice));
  }
  return SPRAYGPU_OK;
}

static int image_launch_flush(Image* im, float4* target, int to_root, int to_host, int clear, cudaStream_t st, int threads) {
  float4* root = nullptr;
  if (to_root && im->nranks > 1) root = (float4*)im->root_mapped[(size_t)im->root];
  float4* host = to_host ? im->h_image_dev : nullptr;
  if (!root && !host && !clear) return SPRAYGPU_OK;
  const size_t max_pix = (size_t)im->max_rows * im->w;
  // ~2 warps of work per 256-thread block pass; enough blocks to cover the SMs, a grid-stride loop does the rest
  int gx = (int)std::min<size_t>((max_pix + 1023) / 1024, (size_t)std::max(1, im->dev->sm_count * 8 / std::max(im->nmy, 1)));
  if (gx < 1) gx = 1;
  image_flush_kernel<<<dim3((unsigned)gx, (unsigned)im->nmy), threads, 0, st>>>(
      target, root, host, im->d_prev_root, im->d_prev_host, im->d_blocks, im->w, clear, im->d_flushed);
  im->dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

int spraygpu_image_flush(spraygpu_image img, int to_root, int to_host, int clear) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_flush: img is NULL");
  Image* im = (Image*)img;
  if (!im->nmy) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  if (im->flush_stream) SG_CUDA(cudaStreamSynchronize(im->flush_stream));   // the bitmaps are shared with pending async flushes
  return image_launch_flush(im, im->d_target, to_root, to_host, clear, im->dev->stream, 256);
}

// Frame-level double buffering: the flush of the frame just traced runs on its own (high-priority) stream with
// 128-thread CTAs of 30 registers — they fit beside the five resident CTAs of the persistent render kernel on every
// SM — so its PCIe / NVLink stores overlap the NEXT frame's trace, which goes to the other render target.  The target
// is always cleared (it is traced into again two frames later).
int spraygpu_image_flush_async(spraygpu_image img, int to_root, int to_host) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_flush_async: img is NULL");
  Image* im = (Image*)img;
  Device* dev = im->dev;
  SG_CUDA(cudaSetDevice(dev->cuda_device));
  if (!im->flush_stream) {
    int lo = 0, hi = 0;
    SG_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    SG_CUDA(cudaStreamCreateWithPriority(&im->flush_stream, cudaStreamNonBlocking, hi));
    SG_CUDA(cudaEventCreateWithFlags(&im->ev_trace, cudaEventDisableTiming));
    for (int k = 0; k < 2; ++k) SG_CUDA(cudaEventCreateWithFlags(&im->ev_flush[k], cudaEventDisableTiming));
    SG_CUDA(cudaMalloc((void**)&im->d_targets[1], im->bytes));
    SG_CUDA(cudaMemsetAsync(im->d_targets[1], 0, im->bytes, dev->stream));
  }
  SG_CUDA(cudaEventRecord(im->ev_trace, dev->stream));
  SG_CUDA(cudaStreamWaitEvent(im->flush_stream, im->ev_trace, 0));
  if (im->nmy) {
    int rc = image_launch_flush(im, im->d_targets[im->cur], to_root, to_host, 1, im->flush_stream, 128);
    if (rc) return rc;
  }
  SG_CUDA(cudaEventRecord(im->ev_flush[im->cur], im->flush_stream));
  im->flushed_once[im->cur] = true;
  im->cur ^= 1;
  im->d_target = im->d_targets[im->cur];
  // the next frame is traced into the other target: its own flush (two frames ago) must have cleared it
  if (im->flushed_once[im->cur]) SG_CUDA(cudaStreamWaitEvent(dev->stream, im->ev_flush[im->cur], 0));
  return SPRAYGPU_OK;
}

// host wait: every flush issued so far has landed (the host image holds the last flushed frame)
int spraygpu_image_wait(spraygpu_image img) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_wait: img is NULL");
  Image* im = (Image*)img;
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  if (im->flush_stream) SG_CUDA(cudaStreamSynchronize(im->flush_stream));
  return SPRAYGPU_OK;
}

int spraygpu_image_flushed_pixels(spraygpu_image img, uint64_t* to_root, uint64_t* to_host) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_flushed_pixels: img is NULL");
  Image* im = (Image*)img;
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  if (im->flush_stream) SG_CUDA(cudaStreamSynchronize(im->flush_stream));
  unsigned long long h[2] = {0, 0};
  SG_CUDA(cudaMemcpyAsync(h, im->d_flushed, sizeof(h), cudaMemcpyDeviceToHost, im->dev->stream));
  SG_CUDA(cudaStreamSynchronize(im->dev->stream));
  if (to_root) *to_root = h[0];
  if (to_host) *to_host = h[1];
  return SPRAYGPU_OK;
}

int spraygpu_image_read(spraygpu_image img, int which, float* out_host) {
  SG_CHECK(img && out_host, SPRAYGPU_ERR_ARG, "image_read: NULL argument");
  Image* im = (Image*)img;
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  const float4* src = im->d_target;
  if (which == 1) src = im->nranks > 1 ? (const float4*)im->root_mapped[(size_t)im->root] : im->d_target;
  SG_CUDA(cudaMemcpyAsync(out_host, src, im->bytes, cudaMemcpyDeviceToHost, im->dev->stream));
  SG_CUDA(cudaStreamSynchronize(im->dev->stream));
  return SPRAYGPU_OK;
}

int spraygpu_image_barrier(spraygpu_image img) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_barrier: img is NULL");
  Image* im = (Image*)img;
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  SG_CUDA(cudaStreamSynchronize(im->dev->stream));
  if (im->flush_stream) SG_CUDA(cudaStreamSynchronize(im->flush_stream));
  if (im->comm) return comm_host_barrier(im->comm);
  return SPRAYGPU_OK;
}

}  // extern "C"
