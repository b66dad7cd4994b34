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
int comm_all_ok(Comm* c, int ok);
int comm_rank(const Comm* c);
int comm_size(const Comm* c);
Device* comm_device(const Comm* c);

struct Image {
  Device* dev = nullptr;
  Comm* comm = nullptr;
  int w = 0, h = 0, rank = 0, nranks = 1, root = 0;
  float4* d_target = nullptr;          // this rank's render target
  // spraygpu_image_flush_async: the host-bound pixels of frame k are COMPACTED on the device's stream (index + value
  // lists, double-buffered) and stored into the host image by a few CTAs on `flush_stream` while frame k + 1 is traced
  uint32_t* d_list_idx[2] = {nullptr, nullptr};
  float4* d_list_val[2] = {nullptr, nullptr};
  unsigned int* d_list_n = nullptr;    // [2] entries in each list
  int cur = 0;
  cudaStream_t flush_stream = nullptr;
  cudaEvent_t ev_compact = nullptr, ev_store[2] = {nullptr, nullptr};
  bool stored_once[2] = {false, false};
  float4* d_root_local = nullptr;      // root: the composite image (peer-mapped into every rank); others: nullptr
  std::vector<void*> root_mapped;      // [nranks] mappings of every rank's shared buffer; [root] is the composite
  float4* h_image = nullptr;           // pinned host image (private cudaHostAlloc, or shared memory + cudaHostRegister)
  float4* h_image_dev = nullptr;       // the address kernels use for it (cudaHostGetDevicePointer)
  std::string shm_name;
  bool shm_owner = false;
  size_t bytes = 0;
  // rows this rank renders: blocks [row_begin[b], row_begin[b+1]) with owner[b] == rank
  std::vector<int2> my_blocks;         // host copy of d_blocks (set_rows compares the old and the new ownership)
  int2* d_blocks = nullptr;            // (first row, row count) of my blocks
  int nmy = 0, max_rows = 0;
  uint32_t* d_prev_root = nullptr;     // 1 bit / pixel: the root composite holds a non-zero value there
  uint32_t* d_prev_host = nullptr;     // 1 bit / pixel: the host image holds a non-zero value there
  unsigned long long* d_flushed = nullptr;   // [0] pixels stored to root, [1] pixels stored to host (cumulative)
};

// One warp handles FLUSH_GROUPS x 32 consecutive pixels of a row block per iteration: the loads of all groups are issued
// before any is used, so a warp keeps FLUSH_GROUPS x 512 B in flight.  That matters for the overlapped flush
// (spraygpu_image_flush_async), whose CTAs squeeze in beside the persistent render kernel — 3 warps per SM: with one
// group per iteration the scan of a 64 MB target was latency-bound at 0.8 ms (r04p), longer than the frame it was
// meant to hide behind.  The two "what the destination holds" bitmaps are read and written one word per group.
#define FLUSH_GROUPS 4
__global__ void __launch_bounds__(256, 4)
    image_flush_kernel(float4* __restrict__ target, float4* __restrict__ root, float4* __restrict__ host, uint32_t* __restrict__ prev_root,
                       uint32_t* __restrict__ prev_host, const int2* __restrict__ blocks, int image_w, int clear,
                       unsigned long long* __restrict__ flushed, uint32_t* __restrict__ list_idx, float4* __restrict__ list_val,
                       unsigned int* __restrict__ list_n) {
  // list_idx != NULL: the host-bound pixels are appended to (list_idx, list_val) instead of being stored to `host`
  // (image_store_kernel does that later); `host` then only says that the host image is a destination
  const int2 blk = blocks[blockIdx.y];
  const size_t npix = (size_t)blk.y * image_w;
  const size_t base = (size_t)blk.x * image_w;            // first pixel of the block: a multiple of 32 (image_w % 32 == 0)
  const int lane = threadIdx.x & 31;
  const size_t warps = ((size_t)gridDim.x * blockDim.x) >> 5;
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  unsigned long long n_root = 0, n_host = 0;
  for (size_t p0 = warp * (32 * FLUSH_GROUPS); p0 < npix; p0 += warps * (32 * FLUSH_GROUPS)) {
    float4 v[FLUSH_GROUPS];
    uint32_t pr = 0u, ph = 0u;                              // lane g holds group g's bitmap words
    const size_t word0 = (base + p0) >> 5;
    if (lane < FLUSH_GROUPS && p0 + 32 * (size_t)lane < npix) {
      if (root) pr = prev_root[word0 + lane];
      if (host) ph = prev_host[word0 + lane];
    }
#pragma unroll
    for (int g = 0; g < FLUSH_GROUPS; ++g) {
      const size_t p = p0 + 32 * (size_t)g + lane;
      v[g] = p < npix ? target[base + p] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int g = 0; g < FLUSH_GROUPS; ++g) {
      const size_t p = p0 + 32 * (size_t)g + lane;
      const bool in = p < npix;
      const bool nz = in && (v[g].x != 0.0f || v[g].y != 0.0f || v[g].z != 0.0f || v[g].w != 0.0f);
      const uint32_t nzm = __ballot_sync(0xffffffffu, nz);
      if (root) {
        const uint32_t pm = __shfl_sync(0xffffffffu, pr, g);
        if (in && ((nzm | pm) >> lane & 1u)) root[base + p] = v[g];
        if (lane == 0 && (nzm | pm)) {
          if (pm != nzm) prev_root[word0 + g] = nzm;
          n_root += (unsigned long long)__popc(nzm | pm);
        }
      }
      if (host) {
        const uint32_t pm = __shfl_sync(0xffffffffu, ph, g);
        const uint32_t m = nzm | pm;
        if (list_idx) {
          uint32_t slot = 0;
          if (lane == 0 && m) slot = atomicAdd(list_n, (unsigned int)__popc(m));
          slot = __shfl_sync(0xffffffffu, slot, 0);
          if (in && (m >> lane & 1u)) {
            const uint32_t e = slot + __popc(m & ((1u << lane) - 1u));
            list_idx[e] = (uint32_t)(base + p);
            list_val[e] = v[g];
          }
        } else if (in && (m >> lane & 1u)) host[base + p] = v[g];
        if (lane == 0 && m) {
          if (pm != nzm) prev_host[word0 + g] = nzm;
          n_host += (unsigned long long)__popc(m);
        }
      }
      if (clear && nz) target[base + p] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  if (flushed && lane == 0) {
    if (n_root) atomicAdd(flushed + 0, n_root);
    if (n_host) atomicAdd(flushed + 1, n_host);
  }
}

// The PCIe half of spraygpu_image_flush_async: a FEW CTAs stream the compact list into the pinned host image while the
// next frame is traced.  Measured (tools/pipe_probe.py, r04r): when every SM carries a CTA that issues zero-copy stores
// to system memory, the persistent render kernel beside it slows from 0.66 to 0.84 ms — PCIe back-pressure blocks the
// SM's path to L2 for everybody on that SM — so the stores must come from few SMs, and the scan of the 64 MB target
// (which needs all of them to be fast) is done separately, without stores, on the device's stream.
__global__ void __launch_bounds__(64) image_store_kernel(const uint32_t* __restrict__ list_idx, const float4* __restrict__ list_val,
                                                         const unsigned int* __restrict__ list_n, float4* __restrict__ host) {
  const unsigned int n = *list_n;
  const unsigned int stride = gridDim.x * blockDim.x;
  unsigned int e = blockIdx.x * blockDim.x + threadIdx.x;
  for (; e + 3 * stride < n; e += 4 * stride) {          // four independent 20-byte fetches in flight per lane
    const uint32_t i0 = list_idx[e], i1 = list_idx[e + stride], i2 = list_idx[e + 2 * stride], i3 = list_idx[e + 3 * stride];
    const float4 v0 = list_val[e], v1 = list_val[e + stride], v2 = list_val[e + 2 * stride], v3 = list_val[e + 3 * stride];
    host[i0] = v0; host[i1] = v1; host[i2] = v2; host[i3] = v3;
  }
  for (; e < n; e += stride) host[list_idx[e]] = list_val[e];
}

// Rows that LEAVE this rank (spraygpu_image_set_rows): the "what the destination holds" bits are per rank, so the new
// owner knows nothing about the non-zero pixels this rank's flushes left at the destinations and would never overwrite
// the ones it renders black.  The old owner therefore zeroes them there and forgets them.  One warp per bitmap word.
__global__ void __launch_bounds__(256) image_release_kernel(float4* __restrict__ root, float4* __restrict__ host,
                                                            uint32_t* __restrict__ prev_root, uint32_t* __restrict__ prev_host,
                                                            const int2* __restrict__ ranges, int image_w) {
  const int2 r = ranges[blockIdx.y];
  const size_t word0 = ((size_t)r.x * image_w) >> 5, words = ((size_t)r.y * image_w) >> 5;
  const int lane = threadIdx.x & 31;
  const size_t warps = ((size_t)gridDim.x * blockDim.x) >> 5;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  for (size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < words; w += warps) {
    const uint32_t mr = root ? prev_root[word0 + w] : 0u, mh = host ? prev_host[word0 + w] : 0u;
    if (mr >> lane & 1u) root[((word0 + w) << 5) + lane] = zero;
    if (mh >> lane & 1u) host[((word0 + w) << 5) + lane] = zero;
    __syncwarp();
    if (lane == 0) {
      if (mr) prev_root[word0 + w] = 0u;
      if (mh) prev_host[word0 + w] = 0u;
    }
  }
}

}  // namespace spraygpu

using namespace spraygpu;

static int image_set_rows_impl(Image* im, int nblocks, const int32_t* row_begin, const int32_t* owner, bool collective);

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
    cudaMemsetAsync(im->d_target, 0, im->bytes, dev->stream);
    cudaMemsetAsync(im->d_prev_root, 0, words * 4, dev->stream);
    cudaMemsetAsync(im->d_prev_host, 0, words * 4, dev->stream);
    cudaMemsetAsync(im->d_flushed, 0, 16, dev->stream);
    // Collective set-up steps end in comm_all_ok(): a rank that failed keeps walking through the remaining rendezvous
    // points instead of bailing out alone, and EVERY rank returns the error — a lone early return would leave the
    // peers holding an image one rank does not have, hanging in the next collective.
    if (comm && im->nranks > 1) {
      // the composite lives on the root; the other ranks allocate a token buffer so that the exchange stays symmetric
      void* local = nullptr;
      const int arc = comm_shared_alloc(comm, im->rank == im->root ? im->bytes : 256, &local, &im->root_mapped, im->root);
      im->d_root_local = (float4*)local;
      const int all = comm_all_ok(comm, arc == SPRAYGPU_OK);
      if (all != 1) {
        if (arc == SPRAYGPU_OK) set_error("image_create: another rank could not set up the peer-mapped composite");
        // everybody is here (comm_all_ok is a rendezvous): drop the mappings now, so that no rank's destroy waits at a
        // barrier the rank that failed never reaches
        comm_shared_free(comm, im->d_root_local, &im->root_mapped);
        im->d_root_local = nullptr;
        rc = arc ? arc : SPRAYGPU_ERR_STATE; break;
      }
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
      // the segment exists before anyone opens it — or nobody goes on
      if (comm && comm_all_ok(comm, im->rank != im->root || fd >= 0) != 1) {
        if (fd >= 0) close(fd);
        set_error("image_create: the root could not create the shared-memory segment %s", im->shm_name.c_str());
        rc = SPRAYGPU_ERR_STATE; break;
      }
      int step = 0;       // 0 ok, else what failed on THIS rank
      if (im->rank != im->root) fd = shm_open(im->shm_name.c_str(), O_RDWR, 0600);
      if (fd < 0) step = 1;
      void* p = MAP_FAILED;
      if (!step) {
        p = mmap(nullptr, im->bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        close(fd);
        if (p == MAP_FAILED) step = 2;
      }
      if (!step) {
        if (im->rank == im->root) memset(p, 0, im->bytes);
        im->h_image = (float4*)p;
        if (cudaHostRegister(p, im->bytes, cudaHostRegisterPortable | cudaHostRegisterMapped) != cudaSuccess) {
          cudaGetLastError();
          munmap(p, im->bytes); im->h_image = nullptr;      // destroy must not unregister what was never registered
          step = 3;
        }
      }
      // zero-filled and mapped everywhere before anyone flushes into it
      const int all = comm ? comm_all_ok(comm, step == 0) : (step == 0);
      if (all != 1) {
        static const char* what[] = {"another rank failed to map the shared host image", "shm_open failed", "mmap failed", "cudaHostRegister failed"};
        set_error("image_create: %s (%s)", what[step], im->shm_name.c_str());
        rc = step == 3 ? SPRAYGPU_ERR_CUDA : SPRAYGPU_ERR_STATE; break;
      }
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
  return image_set_rows_impl(im, 1, rb, nullptr, false);   // nothing flushed yet: nothing to hand over, no rendezvous
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
  if (im->ev_compact) cudaEventDestroy(im->ev_compact);
  for (int k = 0; k < 2; ++k) {
    if (im->ev_store[k]) cudaEventDestroy(im->ev_store[k]);
    if (im->d_list_idx[k]) cudaFree(im->d_list_idx[k]);
    if (im->d_list_val[k]) cudaFree(im->d_list_val[k]);
  }
  if (im->d_list_n) cudaFree(im->d_list_n);
  if (im->d_target) cudaFree(im->d_target);
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
  return image_set_rows_impl((Image*)img, nblocks, row_begin, owner, true);
}

static int image_launch_flush(Image* im, int to_root, int to_host, int clear, cudaStream_t st, int list /* -1: store directly */) {
  float4* root = nullptr;
  if (to_root && im->nranks > 1) root = (float4*)im->root_mapped[(size_t)im->root];
  float4* host = to_host ? im->h_image_dev : nullptr;
  if (!root && !host && !clear) return SPRAYGPU_OK;
  const size_t max_pix = (size_t)im->max_rows * im->w;
  const int threads = 256;
  // enough blocks to cover the SMs, a grid-stride loop does the rest
  int gx = (int)std::min<size_t>((max_pix + (size_t)threads * FLUSH_GROUPS - 1) / ((size_t)threads * FLUSH_GROUPS),
                                 (size_t)std::max(1, im->dev->sm_count * 8 / std::max(im->nmy, 1)));
  if (gx < 1) gx = 1;
  const bool use_list = list >= 0 && host;
  image_flush_kernel<<<dim3((unsigned)gx, (unsigned)im->nmy), threads, 0, st>>>(
      im->d_target, root, host, im->d_prev_root, im->d_prev_host, im->d_blocks, im->w, clear, im->d_flushed,
      use_list ? im->d_list_idx[list] : nullptr, use_list ? im->d_list_val[list] : nullptr, use_list ? im->d_list_n + list : nullptr);
  im->dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

int spraygpu_image_flush(spraygpu_image img, int to_root, int to_host, int clear) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_flush: img is NULL");
  Image* im = (Image*)img;
  if (!im->nmy) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  // a synchronous flush to the host must not overtake stores of earlier asynchronous ones
  if (im->flush_stream && to_host) SG_CUDA(cudaStreamSynchronize(im->flush_stream));
  return image_launch_flush(im, to_root, to_host, clear, im->dev->stream, -1);
}

// The flush OVERLAPPED with the next frame.  On the device's stream, right behind the trace: scan the render target,
// push the root-bound pixels over NVLink as before, APPEND the host-bound pixels (index + value) to a compact device
// list and clear the target — no system-memory store, ~25 us for a 64 MB target.  On the image's own stream: a few
// small CTAs (they fit into the registers the persistent render kernel leaves free) stream that list into the pinned
// host image while the next frame is traced.  Two lists alternate; a list is re-used only after its stores are done.
int spraygpu_image_flush_async(spraygpu_image img, int to_root, int to_host) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_flush_async: img is NULL");
  Image* im = (Image*)img;
  Device* dev = im->dev;
  SG_CUDA(cudaSetDevice(dev->cuda_device));
  if (!to_host) return spraygpu_image_flush(img, to_root, 0, 1);
  if (!im->flush_stream) {
    int least = 0, greatest = 0;
    SG_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    SG_CUDA(cudaStreamCreateWithPriority(&im->flush_stream, cudaStreamNonBlocking, greatest));
    SG_CUDA(cudaEventCreateWithFlags(&im->ev_compact, cudaEventDisableTiming));
    const size_t npix = (size_t)im->w * im->h;
    for (int k = 0; k < 2; ++k) {
      SG_CUDA(cudaEventCreateWithFlags(&im->ev_store[k], cudaEventDisableTiming));
      SG_CUDA(cudaMalloc((void**)&im->d_list_idx[k], npix * sizeof(uint32_t)));
      SG_CUDA(cudaMalloc((void**)&im->d_list_val[k], npix * sizeof(float4)));
    }
    SG_CUDA(cudaMalloc((void**)&im->d_list_n, sizeof(unsigned int) * 2));
  }
  const int k = im->cur;
  if (im->stored_once[k]) SG_CUDA(cudaStreamWaitEvent(dev->stream, im->ev_store[k], 0));   // list k is free again
  SG_CUDA(cudaMemsetAsync(im->d_list_n + k, 0, sizeof(unsigned int), dev->stream));
  if (im->nmy) {
    int rc = image_launch_flush(im, to_root, 1, 1, dev->stream, k);
    if (rc) return rc;
  }
  SG_CUDA(cudaEventRecord(im->ev_compact, dev->stream));
  SG_CUDA(cudaStreamWaitEvent(im->flush_stream, im->ev_compact, 0));
  image_store_kernel<<<16, 64, 0, im->flush_stream>>>(im->d_list_idx[k], im->d_list_val[k], im->d_list_n + k, im->h_image_dev);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  SG_CUDA(cudaEventRecord(im->ev_store[k], im->flush_stream));
  im->stored_once[k] = true;
  im->cur ^= 1;
  return SPRAYGPU_OK;
}

// device-side join: work enqueued on the device's stream after this call starts only when every flush issued so far has
// landed (no host wait) — e.g. a kernel that reads the host image, or an event that closes a timed region
int spraygpu_image_join(spraygpu_image img) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_join: img is NULL");
  Image* im = (Image*)img;
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  for (int k = 0; k < 2; ++k)
    if (im->stored_once[k]) SG_CUDA(cudaStreamWaitEvent(im->dev->stream, im->ev_store[k], 0));
  return SPRAYGPU_OK;
}

// host wait: every flush issued so far has landed (the host image holds the last flushed frame)
int spraygpu_image_wait(spraygpu_image img) {
  SG_CHECK(img, SPRAYGPU_ERR_ARG, "image_wait: img is NULL");
  Image* im = (Image*)img;
  SG_CUDA(cudaSetDevice(im->dev->cuda_device));
  SG_CUDA(cudaStreamSynchronize(im->dev->stream));
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

// set_rows proper.  Rows this rank gives up are released first (image_release_kernel): pending asynchronous stores are
// waited for, the pixels this rank left non-zero at the root composite / the host image are zeroed there, and — with a
// communicator — all ranks meet, so that a new owner's first flush cannot be overtaken by the old owner's zeroes.
static int image_set_rows_impl(Image* im, int nblocks, const int32_t* row_begin, const int32_t* owner, bool collective) {
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
  if (im->flush_stream) SG_CUDA(cudaStreamSynchronize(im->flush_stream));
  // rows that were mine and are not any more
  std::vector<int2> released;
  if (!im->my_blocks.empty()) {
    std::vector<char> keep((size_t)im->h, 0), had((size_t)im->h, 0);
    for (const int2& b : mine) std::fill(keep.begin() + b.x, keep.begin() + b.x + b.y, 1);
    for (const int2& b : im->my_blocks) std::fill(had.begin() + b.x, had.begin() + b.x + b.y, 1);
    for (int y = 0; y < im->h;) {
      if (!(had[(size_t)y] && !keep[(size_t)y])) { ++y; continue; }
      int y1 = y;
      while (y1 < im->h && had[(size_t)y1] && !keep[(size_t)y1]) ++y1;
      released.push_back(make_int2(y, y1 - y));
      y = y1;
    }
  }
  if (!released.empty()) {
    float4* root = im->nranks > 1 ? (float4*)im->root_mapped[(size_t)im->root] : nullptr;
    int2* d_ranges = nullptr;
    SG_CUDA(cudaMalloc((void**)&d_ranges, sizeof(int2) * released.size()));
    cudaError_t e = cudaMemcpy(d_ranges, released.data(), sizeof(int2) * released.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
      image_release_kernel<<<dim3((unsigned)std::max(1, im->dev->sm_count * 4 / (int)released.size()), (unsigned)released.size()), 256, 0,
                             im->dev->stream>>>(root, im->h_image_dev, im->d_prev_root, im->d_prev_host, d_ranges, im->w);
      im->dev->launches++;
      e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaStreamSynchronize(im->dev->stream);
    }
    cudaFree(d_ranges);
    if (e != cudaSuccess) { set_error("image_set_rows: releasing rows failed: %s", cudaGetErrorString(e)); return SPRAYGPU_ERR_CUDA; }
  }
  if (collective && im->comm) { int rc = comm_host_barrier(im->comm); if (rc) return rc; }
  if (im->d_blocks) { cudaFree(im->d_blocks); im->d_blocks = nullptr; }
  im->nmy = (int)mine.size(); im->max_rows = max_rows;
  im->my_blocks = mine;
  if (im->nmy) {
    SG_CUDA(cudaMalloc((void**)&im->d_blocks, sizeof(int2) * mine.size()));
    SG_CUDA(cudaMemcpy(im->d_blocks, mine.data(), sizeof(int2) * mine.size(), cudaMemcpyHostToDevice));
  }
  return SPRAYGPU_OK;
}
