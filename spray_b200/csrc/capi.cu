// capi.cu — the C ABI of libspraygpu (include/spray_gpu.h).  Thin: argument checks, host/device pointer
// staging, and launches.  There is NO CPU fallback anywhere: every entry point that computes needs a
// CUDA device and fails with SPRAYGPU_ERR_CUDA otherwise.
#include <limits.h>
#include <math.h>
#include <stdarg.h>

#include <algorithm>

#include <mutex>

#include "internal.cuh"

namespace spraygpu {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static bool is_device_ptr(const void* p) {
  if (!p) return false;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// Per-call staging of host buffers.  Inputs are copied H2D up front; outputs are copied D2H by finish().
struct Stager {
  Device* dev;
  struct Out { void* host; void* devp; size_t bytes; };
  std::vector<Out> outs;
  size_t in_off = 0, out_off = 0;
  bool any_host = false;
  int rc = SPRAYGPU_OK;

  explicit Stager(Device* d) : dev(d) {}

  // must be called before in()/out(): total bytes of host-side inputs / outputs (upper bounds)
  int reserve(size_t in_bytes, size_t out_bytes) {
    int r = ensure_scratch(dev->stage_in, in_bytes + 4096);
    if (r) return r;
    return ensure_scratch(dev->stage_out, out_bytes + 4096);
  }
  template <typename T> const T* in(const T* p, size_t count) {
    if (!p || !count || is_device_ptr(p)) return p;
    any_host = true;
    in_off = (in_off + 255) & ~(size_t)255;
    char* d = (char*)dev->stage_in.ptr + in_off;
    size_t bytes = count * sizeof(T);
    if (in_off + bytes > dev->stage_in.bytes) { set_error("staging overflow (in)"); rc = SPRAYGPU_ERR_STATE; return nullptr; }
    in_off += bytes;
    cudaError_t e = cudaMemcpyAsync(d, p, bytes, cudaMemcpyHostToDevice, dev->stream);
    if (e != cudaSuccess) { set_error("H2D copy failed: %s", cudaGetErrorString(e)); rc = SPRAYGPU_ERR_CUDA; return nullptr; }
    return (const T*)d;
  }
  // copy_in: the kernel reads-modifies the buffer (merge modes)
  template <typename T> T* out(T* p, size_t count, bool copy_in = false) {
    if (!p || !count || is_device_ptr(p)) return p;
    any_host = true;
    out_off = (out_off + 255) & ~(size_t)255;
    char* d = (char*)dev->stage_out.ptr + out_off;
    size_t bytes = count * sizeof(T);
    if (out_off + bytes > dev->stage_out.bytes) { set_error("staging overflow (out)"); rc = SPRAYGPU_ERR_STATE; return nullptr; }
    out_off += bytes;
    if (copy_in) {
      cudaError_t e = cudaMemcpyAsync(d, p, bytes, cudaMemcpyHostToDevice, dev->stream);
      if (e != cudaSuccess) { set_error("H2D copy failed: %s", cudaGetErrorString(e)); rc = SPRAYGPU_ERR_CUDA; return nullptr; }
    }
    outs.push_back({(void*)p, (void*)d, bytes});
    return (T*)d;
  }
  int finish() {
    if (rc) return rc;
    for (auto& o : outs) SG_CUDA(cudaMemcpyAsync(o.host, o.devp, o.bytes, cudaMemcpyDeviceToHost, dev->stream));
    if (any_host) SG_CUDA(cudaStreamSynchronize(dev->stream));
    return SPRAYGPU_OK;
  }
};

// ---- top-level BVH4 over the resident domains' world boxes (replaces WbvhEmbree::build,
// src/render/wbvh_embree.cc:58-111, an Embree user-geometry BVH over the same boxes) ----
static void tlas_bounds(const std::vector<DomainEntry>& doms, const int* ids, int n, float lo[3], float hi[3]) {
  for (int a = 0; a < 3; ++a) { lo[a] = INFINITY; hi[a] = -INFINITY; }
  for (int i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) {
      // boxes may be inverted by a mirroring transform (scene_loader.cc:352-357 keeps min/max corners as given)
      const float b0 = doms[ids[i]].box[a], b1 = doms[ids[i]].box[3 + a];
      lo[a] = fminf(lo[a], fminf(b0, b1)); hi[a] = fmaxf(hi[a], fmaxf(b0, b1));
    }
}
static void tlas_split(const std::vector<DomainEntry>& doms, int* ids, int n, int* nleft) {
  // median split along the axis with the widest centroid spread
  float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) {
      const float c = 0.5f * (doms[ids[i]].box[a] + doms[ids[i]].box[3 + a]);
      clo[a] = fminf(clo[a], c); chi[a] = fmaxf(chi[a], c);
    }
  int ax = 0;
  if (chi[1] - clo[1] > chi[ax] - clo[ax]) ax = 1;
  if (chi[2] - clo[2] > chi[ax] - clo[ax]) ax = 2;
  std::stable_sort(ids, ids + n, [&](int x, int y) {
    return doms[x].box[ax] + doms[x].box[3 + ax] < doms[y].box[ax] + doms[y].box[3 + ax];
  });
  *nleft = n / 2;
}
static int tlas_build(Device* dev, int* ids, int n, int depth, int* max_depth) {
  const int me = (int)dev->h_tlas.size();
  if (depth > *max_depth) *max_depth = depth;
  dev->h_tlas.push_back(TlasNode());
  // up to four groups: halves of halves
  int start[4], cnt[4], ng = 0;
  if (n <= 4) {
    for (int i = 0; i < n; ++i) { start[ng] = i; cnt[ng] = 1; ++ng; }
  } else {
    int nl; tlas_split(dev->h_domains, ids, n, &nl);
    int half[2][2] = {{0, nl}, {nl, n - nl}};
    for (int h = 0; h < 2; ++h) {
      if (half[h][1] >= 2) {
        int q; tlas_split(dev->h_domains, ids + half[h][0], half[h][1], &q);
        start[ng] = half[h][0]; cnt[ng] = q; ++ng;
        start[ng] = half[h][0] + q; cnt[ng] = half[h][1] - q; ++ng;
      } else {
        start[ng] = half[h][0]; cnt[ng] = half[h][1]; ++ng;
      }
    }
  }
  TlasNode node;
  for (int c = 0; c < 4; ++c) {
    for (int a = 0; a < 3; ++a) { node.lo[a][c] = INFINITY; node.hi[a][c] = -INFINITY; }
    node.child[c] = INT32_MIN;
  }
  for (int g = 0; g < ng; ++g) {
    float lo[3], hi[3];
    tlas_bounds(dev->h_domains, ids + start[g], cnt[g], lo, hi);
    for (int a = 0; a < 3; ++a) { node.lo[a][g] = lo[a]; node.hi[a][g] = hi[a]; }
    node.child[g] = cnt[g] == 1 ? ~ids[start[g]] : tlas_build(dev, ids + start[g], cnt[g], depth + 1, max_depth);
  }
  dev->h_tlas[me] = node;
  return me;
}

int upload_domain_table(Device* dev) {
  // every kernel that reads the table comes through here: builds that were only enqueued (spraygpu_scene_commit_async)
  // are waited for and checked now
  for (int i = 0; i < dev->ndomains; ++i)
    if (dev->bound[i] && dev->bound[i]->pending) {
      int frc = scene_finalize(dev->bound[i]);
      if (frc) return frc;
    }
  if (!dev->table_dirty) return SPRAYGPU_OK;
  SG_CHECK(dev->ndomains > 0, SPRAYGPU_ERR_STATE, "no domain table: call spraygpu_domains_set first");
  {
    std::vector<int> ids;
    for (int i = 0; i < dev->ndomains; ++i)
      if (dev->h_domains[i].resident && (dev->active.empty() || dev->active[(size_t)i])) ids.push_back(i);
    dev->h_tlas.clear();
    int tlas_depth = 0;
    if (!ids.empty()) tlas_build(dev, ids.data(), (int)ids.size(), 1, &tlas_depth);
    dev->tlas_n = (int)dev->h_tlas.size();
    // TlasSource (traverse.cuh) pops one entry and pushes <= 4 per level: worst-case stack = 3 per level + 1.
    // <= 512 domains split 4 ways per level => <= 5 levels => <= 16 of the 24 entries; checked, not assumed.
    SG_CHECK(3 * tlas_depth + 1 <= 24 /* SPRAY_TLAS_STACK */, SPRAYGPU_ERR_LIMIT, "top-level BVH: depth %d exceeds the traversal stack", tlas_depth);
    SG_CHECK(dev->tlas_n <= SPRAY_TLAS_MAX_NODES, SPRAYGPU_ERR_LIMIT, "top-level BVH: %d nodes", dev->tlas_n);
    if (dev->tlas_n)
      SG_CUDA(cudaMemcpyAsync(dev->d_tlas, dev->h_tlas.data(), sizeof(TlasNode) * dev->tlas_n, cudaMemcpyHostToDevice,
                              dev->stream));
  }
  if (dev->active.empty()) {
    SG_CUDA(cudaMemcpyAsync(dev->d_domains, dev->h_domains.data(), sizeof(DomainEntry) * dev->ndomains,
                            cudaMemcpyHostToDevice, dev->stream));
  } else {
    std::vector<DomainEntry> masked(dev->h_domains);
    for (int i = 0; i < dev->ndomains; ++i)
      if (!dev->active[(size_t)i]) masked[(size_t)i].resident = 0;
    SG_CUDA(cudaMemcpyAsync(dev->d_domains, masked.data(), sizeof(DomainEntry) * dev->ndomains, cudaMemcpyHostToDevice,
                            dev->stream));
  }
  // h_domains is pageable: the runtime stages it before returning, so later host edits are safe
  dev->table_dirty = false;
  return SPRAYGPU_OK;
}

}  // namespace spraygpu

using namespace spraygpu;

#define DEV(d) ((Device*)(d))
#define SCN(s) ((Scene*)(s))

extern "C" {

const char* spraygpu_last_error(void) { return g_err; }
const char* spraygpu_version(void) { return "spraygpu 0.1 (sm_100a)"; }

int spraygpu_device_create(int cuda_device, spraygpu_device* out) {
  SG_CHECK(out, SPRAYGPU_ERR_ARG, "device_create: out is NULL");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    set_error("no CUDA device available (%s); libspraygpu has no CPU fallback",
              e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    cudaGetLastError();
    return SPRAYGPU_ERR_CUDA;
  }
  SG_CHECK(cuda_device >= 0 && cuda_device < count, SPRAYGPU_ERR_ARG, "device_create: bad device %d", cuda_device);
  SG_CUDA(cudaSetDevice(cuda_device));
  Device* d = new Device();
  d->cuda_device = cuda_device;
  cudaDeviceProp prop;
  SG_CUDA(cudaGetDeviceProperties(&prop, cuda_device));
  d->sm_count = prop.multiProcessorCount;
  SG_CUDA(cudaMalloc(&d->d_domains, sizeof(DomainEntry) * SPRAYGPU_MAX_DOMAINS));
  SG_CUDA(cudaMalloc(&d->d_tlas, sizeof(TlasNode) * SPRAY_TLAS_MAX_NODES));
  // 8 render counters + 8 words of scratch that is NOT touched by render launches (set_mesh's max-index reduction
  // runs in the middle of an out-of-core frame, between two counted launches)
  SG_CUDA(cudaMalloc(&d->d_counters, sizeof(unsigned long long) * 16));
  SG_CUDA(cudaMemset(d->d_counters, 0, sizeof(unsigned long long) * 16));
  SG_CUDA(cudaMallocHost(&d->pinned, 4096));
  d->pinned_bytes = 4096;
  SG_CUDA(cudaStreamCreateWithFlags(&d->copy_stream, cudaStreamNonBlocking));
  SG_CUDA(cudaStreamCreateWithFlags(&d->load_stream, cudaStreamNonBlocking));
  SG_CUDA(cudaStreamCreateWithFlags(&d->upload_stream, cudaStreamNonBlocking));
  for (int i = 0; i < 2; ++i) {
    SG_CUDA(cudaEventCreateWithFlags(&d->ev_render[i], cudaEventDisableTiming));
    SG_CUDA(cudaEventCreateWithFlags(&d->ev_copy[i], cudaEventDisableTiming));
  }
  *out = (spraygpu_device)d;
  return SPRAYGPU_OK;
}

int spraygpu_device_destroy(spraygpu_device dev) {
  if (!dev) return SPRAYGPU_OK;
  Device* d = DEV(dev);
  cudaSetDevice(d->cuda_device);
  cudaStreamSynchronize(d->stream);
  if (d->copy_stream) { cudaStreamSynchronize(d->copy_stream); cudaStreamDestroy(d->copy_stream); }
  if (d->load_stream) { cudaStreamSynchronize(d->load_stream); cudaStreamDestroy(d->load_stream); }
  if (d->upload_stream) { cudaStreamSynchronize(d->upload_stream); cudaStreamDestroy(d->upload_stream); }
  for (int k = 0; k < SPRAY_BUILD_POOL; ++k) {
    if (d->pool_stream[k]) { cudaStreamSynchronize(d->pool_stream[k]); cudaStreamDestroy(d->pool_stream[k]); }
    if (d->pool_scratch[k].ptr) cudaFree(d->pool_scratch[k].ptr);
  }
  if (d->pool_order) cudaEventDestroy(d->pool_order);
  for (int i = 0; i < 2; ++i) {
    if (d->ev_render[i]) cudaEventDestroy(d->ev_render[i]);
    if (d->ev_copy[i]) cudaEventDestroy(d->ev_copy[i]);
    if (d->fb_stage[i].ptr) cudaFree(d->fb_stage[i].ptr);
  }
  if (d->scratch.ptr) cudaFree(d->scratch.ptr);
  if (d->stage_in.ptr) cudaFree(d->stage_in.ptr);
  if (d->stage_out.ptr) cudaFree(d->stage_out.ptr);
  if (d->wave.ptr) cudaFree(d->wave.ptr);
  if (d->pinned) cudaFreeHost(d->pinned);
  cudaFree(d->d_domains);
  cudaFree(d->d_tlas);
  cudaFree(d->d_counters);
  delete d;
  return SPRAYGPU_OK;
}

int spraygpu_device_set_stream(spraygpu_device dev, void* cuda_stream) {
  SG_CHECK(dev, SPRAYGPU_ERR_ARG, "set_stream: dev is NULL");
  DEV(dev)->stream = (cudaStream_t)cuda_stream;
  return SPRAYGPU_OK;
}

int spraygpu_device_synchronize(spraygpu_device dev) {
  SG_CHECK(dev, SPRAYGPU_ERR_ARG, "synchronize: dev is NULL");
  SG_CUDA(cudaStreamSynchronize(DEV(dev)->stream));
  if (DEV(dev)->copy_stream) SG_CUDA(cudaStreamSynchronize(DEV(dev)->copy_stream));
  return SPRAYGPU_OK;
}

uint64_t spraygpu_device_launch_count(spraygpu_device dev) { return dev ? DEV(dev)->launches : 0; }

// ---------------------------------------------------------------------------------------------
int spraygpu_scene_create(spraygpu_device dev, spraygpu_scene* out) {
  SG_CHECK(dev && out, SPRAYGPU_ERR_ARG, "scene_create: NULL argument");
  Scene* s = new Scene();
  s->dev = DEV(dev);
  *out = (spraygpu_scene)s;
  return SPRAYGPU_OK;
}

static void scene_free_mesh(Scene* s) {
  if (s->d_verts_xf) cudaFree(s->d_verts_xf);
  if (s->d_verts_obj) cudaFree(s->d_verts_obj);
  if (s->d_faces) cudaFree(s->d_faces);
  if (s->d_rgb) cudaFree(s->d_rgb);
  if (s->d_normals) cudaFree(s->d_normals);
  if (s->d_nodes) cudaFree(s->d_nodes);
  if (s->d_tris) cudaFree(s->d_tris);
  if (s->d_attr) cudaFree(s->d_attr);
  s->d_verts = s->d_verts_obj = s->d_verts_xf = nullptr; s->d_faces = s->d_rgb = nullptr; s->d_normals = nullptr;
  s->d_nodes = nullptr; s->d_tris = s->d_attr = nullptr;
  s->cap_verts_obj = s->cap_faces = s->cap_rgb = s->cap_verts_xf = s->cap_normals = s->cap_tris = s->cap_attr = s->cap_nodes = 0;
  s->committed = false; s->n_nodes = 0;
}

int spraygpu_scene_destroy(spraygpu_scene scene) {
  if (!scene) return SPRAYGPU_OK;
  Scene* s = SCN(scene);
  Device* d = s->dev;
  cudaSetDevice(d->cuda_device);
  cudaStreamSynchronize(d->stream);
  if (d->load_stream) cudaStreamSynchronize(d->load_stream);
  if (s->pending && s->ev1) cudaEventSynchronize(s->ev1);     // a build on a pool stream
  s->pending = false;
  if (s->ev0) { cudaEventDestroy(s->ev0); cudaEventDestroy(s->ev1); }
  if (s->ev_upload) cudaEventDestroy(s->ev_upload);
  if (s->h_result) cudaFreeHost(s->h_result);
  if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
  for (int i = 0; i < d->ndomains; ++i)
    if (d->bound[i] == s) {
      d->bound[i] = nullptr;
      d->h_domains[i].resident = 0;
      d->h_domains[i].nodes = nullptr; d->h_domains[i].tris = nullptr; d->h_domains[i].attr = nullptr;
      d->h_domains[i].verts = nullptr; d->h_domains[i].faces = nullptr;
      d->table_dirty = true;
    }
  scene_free_mesh(s);
  delete s;
  return SPRAYGPU_OK;
}

int spraygpu_scene_set_mesh(spraygpu_scene scene, const float* xyz, size_t nv, const uint32_t* idx, size_t nf,
                            const uint32_t* rgb) {
  SG_CHECK(scene && xyz && idx, SPRAYGPU_ERR_ARG, "set_mesh: NULL argument");
  SG_CHECK(nv > 0 && nf > 0, SPRAYGPU_ERR_ARG, "set_mesh: empty mesh (nv=%zu, nf=%zu)", nv, nf);
  SG_CHECK(nf <= SPRAYGPU_MAX_PRIMS, SPRAYGPU_ERR_LIMIT, "set_mesh: %zu faces exceed the 23-bit primID field", nf);
  SG_CHECK(nv <= 0xFFFFFFFFull, SPRAYGPU_ERR_LIMIT, "set_mesh: too many vertices");
  Scene* s = SCN(scene);
  Device* d = s->dev;
  SG_CUDA(cudaSetDevice(d->cuda_device));
  // The block's buffers are about to be overwritten.  Normally nothing may still read them: wait for the device's
  // stream.  While the out-of-core tile prefetches (ooc.cu) the upload goes to the load stream, which that tile has
  // ordered behind the last trace that used this block.
  cudaStream_t st = upload_stream(d);
  if (!d->loading_async) SG_CUDA(cudaStreamSynchronize(d->stream));
  if (s->pending) { SG_CUDA(cudaEventSynchronize(s->ev1)); s->pending = false; }
  // a re-used cache block keeps its allocations when the new mesh fits (grow_buffer)
  s->committed = false; s->n_nodes = 0; s->d_verts = nullptr;
  s->nv = nv; s->nf = nf;
  int grc;
  if ((grc = grow_buffer(&s->d_verts_obj, &s->cap_verts_obj, sizeof(float) * 3 * nv))) return grc;
  if ((grc = grow_buffer(&s->d_faces, &s->cap_faces, sizeof(uint32_t) * 3 * nf))) return grc;
  SG_CUDA(cudaMemcpyAsync(s->d_verts_obj, xyz, sizeof(float) * 3 * nv, cudaMemcpyDefault, st));
  SG_CUDA(cudaMemcpyAsync(s->d_faces, idx, sizeof(uint32_t) * 3 * nf, cudaMemcpyDefault, st));
  if (rgb) {
    if ((grc = grow_buffer(&s->d_rgb, &s->cap_rgb, sizeof(uint32_t) * nv))) return grc;
    SG_CUDA(cudaMemcpyAsync(s->d_rgb, rgb, sizeof(uint32_t) * nv, cudaMemcpyDefault, st));
  } else if (s->d_rgb) {
    cudaFree(s->d_rgb); s->d_rgb = nullptr; s->cap_rgb = 0;
  }
  // every index must address a vertex (the reference trusts the PLY, ply_loader.cc:243-256; a GPU build would turn a
  // bad index into out-of-bounds device writes).  d_counters[8] is scratch no render launch uses.
  uint32_t* d_max = (uint32_t*)(d->d_counters + 8);
  int vrc = launch_max_index(d, s->d_faces, 3 * nf, d_max);
  if (vrc) return vrc;
  uint32_t* h_max = (uint32_t*)d->pinned;
  SG_CUDA(cudaMemcpyAsync(h_max, d_max, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  if (!s->ev_upload) SG_CUDA(cudaEventCreateWithFlags(&s->ev_upload, cudaEventDisableTiming));
  SG_CUDA(cudaEventRecord(s->ev_upload, st));
  SG_CUDA(cudaStreamSynchronize(st));  // host arrays may be freed after return
  if ((size_t)*h_max >= nv) {
    const uint32_t bad = *h_max;
    s->nv = s->nf = 0;   // the mesh is unusable: commit reports "empty mesh"
    set_error("set_mesh: face index %u out of range (nv = %zu)", bad, nv);
    return SPRAYGPU_ERR_ARG;
  }
  return SPRAYGPU_OK;
}

int spraygpu_scene_set_transform(spraygpu_scene scene, const float* m) {
  SG_CHECK(scene, SPRAYGPU_ERR_ARG, "set_transform: scene is NULL");
  Scene* s = SCN(scene);
  s->has_xf = m != nullptr;
  if (m) memcpy(s->xf, m, sizeof(float) * 16);
  s->committed = false;
  return SPRAYGPU_OK;
}

int spraygpu_scene_commit(spraygpu_scene scene) {
  SG_CHECK(scene, SPRAYGPU_ERR_ARG, "commit: scene is NULL");
  Scene* s = SCN(scene);
  SG_CHECK(s->d_verts_obj && s->d_faces, SPRAYGPU_ERR_STATE, "commit: no mesh set");
  SG_CUDA(cudaSetDevice(s->dev->cuda_device));
  int rc = build_scene(s);
  if (rc) return rc;
  // inside the out-of-core tile's prefetch (ooc.cu) a commit is enqueue-only: the tile waits for the group's builds
  // before it traces the group, and meanwhile the next domain's mesh crosses PCIe
  if (!s->dev->loading_async) {
    rc = scene_finalize(s);
    if (rc) return rc;
  }
  // refresh table entries that point at this scene
  Device* d = s->dev;
  for (int i = 0; i < d->ndomains; ++i)
    if (d->bound[i] == s) spraygpu_domain_bind((spraygpu_device)d, i, scene);
  return SPRAYGPU_OK;
}

int spraygpu_scene_commit_async(spraygpu_scene scene) {
  SG_CHECK(scene, SPRAYGPU_ERR_ARG, "commit_async: scene is NULL");
  Scene* s = SCN(scene);
  SG_CHECK(s->d_verts_obj && s->d_faces, SPRAYGPU_ERR_STATE, "commit_async: no mesh set");
  SG_CUDA(cudaSetDevice(s->dev->cuda_device));
  int rc = build_scene(s);
  if (rc) return rc;
  Device* d = s->dev;
  for (int i = 0; i < d->ndomains; ++i)
    if (d->bound[i] == s) spraygpu_domain_bind((spraygpu_device)d, i, scene);
  return SPRAYGPU_OK;
}

int spraygpu_scene_get_info(spraygpu_scene scene, spraygpu_scene_info* out) {
  SG_CHECK(scene && out, SPRAYGPU_ERR_ARG, "get_info: NULL argument");
  Scene* s = SCN(scene);
  SG_CHECK(s->committed, SPRAYGPU_ERR_STATE, "get_info: scene not committed");
  { int frc = scene_finalize(s); if (frc) return frc; }
  out->num_vertices = s->nv; out->num_faces = s->nf; out->num_nodes = s->n_nodes;
  out->node_bytes = (uint64_t)s->n_nodes * 80; out->tri_bytes = (uint64_t)s->nf * 48; out->attr_bytes = (uint64_t)s->nf * 48;
  memcpy(out->bounds, s->bounds, sizeof(out->bounds));
  out->build_ms = s->build_ms;
  out->bvh_depth = s->bvh_depth; out->reserved = 0;
  return SPRAYGPU_OK;
}

static int copy_out(Device* d, void* dst, const void* src, size_t bytes) {
  SG_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, d->stream));
  if (!is_device_ptr(dst)) SG_CUDA(cudaStreamSynchronize(d->stream));
  return SPRAYGPU_OK;
}
int spraygpu_scene_get_normals(spraygpu_scene scene, float* out) {
  SG_CHECK(scene && out, SPRAYGPU_ERR_ARG, "get_normals: NULL argument");
  Scene* s = SCN(scene);
  SG_CHECK(s->committed, SPRAYGPU_ERR_STATE, "get_normals: scene not committed");
  { int frc = scene_finalize(s); if (frc) return frc; }
  return copy_out(s->dev, out, s->d_normals, sizeof(float) * 3 * s->nv);
}
int spraygpu_scene_get_bvh(spraygpu_scene scene, void* nodes, void* tris) {
  SG_CHECK(scene, SPRAYGPU_ERR_ARG, "get_bvh: scene is NULL");
  Scene* s = SCN(scene);
  SG_CHECK(s->committed, SPRAYGPU_ERR_STATE, "get_bvh: scene not committed");
  { int frc = scene_finalize(s); if (frc) return frc; }
  int rc = SPRAYGPU_OK;
  if (nodes) rc = copy_out(s->dev, nodes, s->d_nodes, (size_t)s->n_nodes * 80);
  if (!rc && tris) rc = copy_out(s->dev, tris, s->d_tris, s->nf * 48);
  return rc;
}
int spraygpu_scene_get_vertices(spraygpu_scene scene, float* out) {
  SG_CHECK(scene && out, SPRAYGPU_ERR_ARG, "get_vertices: NULL argument");
  Scene* s = SCN(scene);
  SG_CHECK(s->committed, SPRAYGPU_ERR_STATE, "get_vertices: scene not committed");
  { int frc = scene_finalize(s); if (frc) return frc; }
  return copy_out(s->dev, out, s->d_verts, sizeof(float) * 3 * s->nv);
}

int spraygpu_intersect(spraygpu_scene scene, int domain_id, const spraygpu_ray* rays, size_t n, spraygpu_key* keys,
                       float* uv) {
  SG_CHECK(scene && (n == 0 || (rays && keys)), SPRAYGPU_ERR_ARG, "intersect: NULL argument");
  Scene* s = SCN(scene);
  SG_CHECK(s->committed, SPRAYGPU_ERR_STATE, "intersect: scene not committed");
  { int frc = scene_finalize(s); if (frc) return frc; }
  SG_CHECK(domain_id >= 0 && domain_id < SPRAYGPU_MAX_DOMAINS, SPRAYGPU_ERR_ARG, "intersect: bad domain id %d", domain_id);
  if (!n) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(s->dev->cuda_device));
  Stager st(s->dev);
  int rc = st.reserve(n * sizeof(spraygpu_ray), n * 16 + 512);
  if (rc) return rc;
  const spraygpu_ray* dr = st.in(rays, n);
  uint64_t* dk = st.out(keys, n);
  float* duv = st.out(uv, 2 * n);
  if (st.rc) return st.rc;
  rc = launch_intersect_scene(s, domain_id, dr, n, dk, duv);
  if (rc) return rc;
  return st.finish();
}

int spraygpu_occluded(spraygpu_scene scene, const spraygpu_ray* rays, size_t n, uint8_t* occ) {
  SG_CHECK(scene && (n == 0 || (rays && occ)), SPRAYGPU_ERR_ARG, "occluded: NULL argument");
  Scene* s = SCN(scene);
  SG_CHECK(s->committed, SPRAYGPU_ERR_STATE, "occluded: scene not committed");
  { int frc = scene_finalize(s); if (frc) return frc; }
  if (!n) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(s->dev->cuda_device));
  Stager st(s->dev);
  int rc = st.reserve(n * sizeof(spraygpu_ray), n + 512);
  if (rc) return rc;
  const spraygpu_ray* dr = st.in(rays, n);
  uint8_t* docc = st.out(occ, n);
  if (st.rc) return st.rc;
  rc = launch_occluded_scene(s, dr, n, docc);
  if (rc) return rc;
  return st.finish();
}

// ---------------------------------------------------------------------------------------------
int spraygpu_domains_set(spraygpu_device dev, int nd, const float* aabb6, const spraygpu_material* mtl) {
  SG_CHECK(dev && aabb6, SPRAYGPU_ERR_ARG, "domains_set: NULL argument");
  SG_CHECK(nd > 0 && nd <= SPRAYGPU_MAX_DOMAINS, SPRAYGPU_ERR_LIMIT, "domains_set: %d domains (limit %d: 9 key bits)", nd,
           SPRAYGPU_MAX_DOMAINS);
  Device* d = DEV(dev);
  d->ndomains = nd;
  d->h_domains.assign(nd, DomainEntry());
  d->bound.assign(nd, nullptr);
  for (int i = 0; i < nd; ++i) {
    DomainEntry& e = d->h_domains[i];
    memset(&e, 0, sizeof(e));
    memcpy(e.box, aabb6 + 6 * i, sizeof(float) * 6);
    e.mtl_type = mtl ? mtl[i].type : SPRAYGPU_MTL_DIFFUSE;
    for (int k = 0; k < 3; ++k) e.mtl[k] = mtl ? mtl[i].params[k] : 1.0f;
  }
  d->table_dirty = true;
  return SPRAYGPU_OK;
}

int spraygpu_domain_bind(spraygpu_device dev, int id, spraygpu_scene scene) {
  SG_CHECK(dev, SPRAYGPU_ERR_ARG, "domain_bind: dev is NULL");
  Device* d = DEV(dev);
  SG_CHECK(id >= 0 && id < d->ndomains, SPRAYGPU_ERR_ARG, "domain_bind: bad domain id %d (table has %d)", id, d->ndomains);
  DomainEntry& e = d->h_domains[id];
  Scene* s = SCN(scene);
  // out-of-core prefetch (ooc.cu): a kernel is reading this domain right now; refusing the eviction here keeps the
  // cache from overwriting its block (the cache is smaller than the max_resident the tile was given)
  SG_CHECK(!(d->loading_async && (size_t)id < d->tracing.size() && d->tracing[(size_t)id] && d->bound[id] != s), SPRAYGPU_ERR_STATE,
           "domain_bind: domain %d is being traced and cannot be evicted now (cache smaller than max_resident?)", id);
  if (s) {
    SG_CHECK(s->committed, SPRAYGPU_ERR_STATE, "domain_bind: scene not committed");
    SG_CHECK(s->dev == d, SPRAYGPU_ERR_ARG, "domain_bind: scene belongs to another device");
    e.resident = 1; e.nprims = (uint32_t)s->nf;
    e.nodes = s->d_nodes; e.tris = s->d_tris; e.attr = s->d_attr; e.verts = s->d_verts; e.faces = s->d_faces;
  } else {
    e.resident = 0; e.nprims = 0;
    e.nodes = nullptr; e.tris = nullptr; e.attr = nullptr; e.verts = nullptr; e.faces = nullptr;
  }
  d->bound[id] = s;
  d->table_dirty = true;
  return SPRAYGPU_OK;
}

int spraygpu_domain_lists(spraygpu_device dev, const spraygpu_ray* rays, size_t n, int cap, int32_t* counts,
                          int32_t* ids, float* tmins) {
  SG_CHECK(dev && (n == 0 || (rays && counts && ids && tmins)) && cap > 0, SPRAYGPU_ERR_ARG, "domain_lists: bad argument");
  Device* d = DEV(dev);
  if (!n) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(d->cuda_device));
  int rc = upload_domain_table(d);
  if (rc) return rc;
  Stager st(d);
  rc = st.reserve(n * sizeof(spraygpu_ray), n * (4 + 8 * (size_t)cap) + 1024);
  if (rc) return rc;
  const spraygpu_ray* dr = st.in(rays, n);
  int32_t* dc = st.out(counts, n);
  int32_t* di = st.out(ids, n * cap);
  float* dt = st.out(tmins, n * cap);
  if (st.rc) return st.rc;
  rc = launch_domain_lists(d, dr, n, cap, dc, di, dt);
  if (rc) return rc;
  return st.finish();
}

int spraygpu_intersect_domains(spraygpu_device dev, const spraygpu_ray* rays, size_t n, spraygpu_key* keys, float* uv,
                               int merge) {
  SG_CHECK(dev && (n == 0 || (rays && keys)), SPRAYGPU_ERR_ARG, "intersect_domains: NULL argument");
  Device* d = DEV(dev);
  if (!n) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(d->cuda_device));
  int rc = upload_domain_table(d);
  if (rc) return rc;
  Stager st(d);
  rc = st.reserve(n * sizeof(spraygpu_ray), n * 16 + 1024);
  if (rc) return rc;
  const spraygpu_ray* dr = st.in(rays, n);
  uint64_t* dk = st.out(keys, n, merge != 0);
  float* duv = st.out(uv, 2 * n, merge != 0);
  if (st.rc) return st.rc;
  rc = launch_intersect_domains(d, dr, n, dk, duv, merge);
  if (rc) return rc;
  return st.finish();
}

int spraygpu_occluded_domains(spraygpu_device dev, const spraygpu_ray* rays, size_t n, uint8_t* occ, int merge) {
  SG_CHECK(dev && (n == 0 || (rays && occ)), SPRAYGPU_ERR_ARG, "occluded_domains: NULL argument");
  Device* d = DEV(dev);
  if (!n) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(d->cuda_device));
  int rc = upload_domain_table(d);
  if (rc) return rc;
  Stager st(d);
  rc = st.reserve(n * sizeof(spraygpu_ray), n + 1024);
  if (rc) return rc;
  const spraygpu_ray* dr = st.in(rays, n);
  uint8_t* docc = st.out(occ, n, merge != 0);
  if (st.rc) return st.rc;
  rc = launch_occluded_domains(d, dr, n, docc, merge);
  if (rc) return rc;
  return st.finish();
}

int spraygpu_hit_attributes(spraygpu_device dev, const spraygpu_ray* rays, const spraygpu_key* keys, size_t n,
                            uint32_t* color, float* ns3) {
  SG_CHECK(dev && (n == 0 || (rays && keys && color && ns3)), SPRAYGPU_ERR_ARG, "hit_attributes: NULL argument");
  Device* d = DEV(dev);
  if (!n) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(d->cuda_device));
  int rc = upload_domain_table(d);
  if (rc) return rc;
  Stager st(d);
  rc = st.reserve(n * (sizeof(spraygpu_ray) + 8) + 1024, n * 16 + 1024);
  if (rc) return rc;
  const spraygpu_ray* dr = st.in(rays, n);
  const uint64_t* dk = st.in(keys, n);
  uint32_t* dc = st.out(color, n);
  float* dn = st.out(ns3, 3 * n);
  if (st.rc) return st.rc;
  rc = launch_hit_attributes(d, dr, dk, n, dc, dn);
  if (rc) return rc;
  return st.finish();
}

static int gen_eye_rays_impl(spraygpu_device dev, const spraygpu_camera* cam, spraygpu_tile tile, int spp, int eye_seed,
                             spraygpu_ray* rays) {
  SG_CHECK(dev && cam && rays && spp > 0, SPRAYGPU_ERR_ARG, "gen_eye_rays: bad argument");
  Device* d = DEV(dev);
  const size_t n = (size_t)tile.w * tile.h * spp;
  if (!n) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(d->cuda_device));
  Stager st(d);
  int rc = st.reserve(0, n * sizeof(spraygpu_ray) + 512);
  if (rc) return rc;
  spraygpu_ray* dr = st.out(rays, n);
  if (st.rc) return st.rc;
  rc = launch_eye_rays(d, cam, tile, spp, eye_seed, dr);
  if (rc) return rc;
  return st.finish();
}
int spraygpu_gen_eye_rays(spraygpu_device dev, const spraygpu_camera* cam, spraygpu_tile tile, int spp, spraygpu_ray* rays) {
  return gen_eye_rays_impl(dev, cam, tile, spp, SPRAYGPU_EYE_SEED_OOC, rays);
}
int spraygpu_gen_eye_rays_insitu(spraygpu_device dev, const spraygpu_camera* cam, spraygpu_tile tile, int spp,
                                 spraygpu_ray* rays) {
  return gen_eye_rays_impl(dev, cam, tile, spp, SPRAYGPU_EYE_SEED_INSITU, rays);
}

static int trace_tile_impl(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                           spraygpu_tile tile, float* fb, spraygpu_stats* stats, bool counted, bool async) {
  SG_CHECK(dev && cam && shade && fb, SPRAYGPU_ERR_ARG, "trace_tile: NULL argument");
  SG_CHECK(shade->pixel_samples >= 1 && shade->bounces >= 1, SPRAYGPU_ERR_ARG, "trace_tile: pixel_samples and bounces must be >= 1");
  SG_CHECK(shade->nlights >= 0 && shade->nlights <= SPRAYGPU_MAX_LIGHTS, SPRAYGPU_ERR_LIMIT, "trace_tile: %d lights (limit %d)",
           shade->nlights, SPRAYGPU_MAX_LIGHTS);
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_AO || shade->nlights > 0, SPRAYGPU_ERR_ARG,
           "trace_tile: the PT shader needs at least one light (ooc_tracer.inl:68)");
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_PT || shade->ao_samples > 0, SPRAYGPU_ERR_ARG, "trace_tile: ao_samples must be > 0");
  SG_CHECK(tile.x >= 0 && tile.y >= 0 && tile.w >= 0 && tile.h >= 0 && tile.x + tile.w <= cam->image_w &&
               tile.y + tile.h <= cam->image_h,
           SPRAYGPU_ERR_ARG, "trace_tile: tile (%d,%d,%d,%d) outside the %dx%d image", tile.x, tile.y, tile.w, tile.h,
           cam->image_w, cam->image_h);
  Device* d = DEV(dev);
  SG_CUDA(cudaSetDevice(d->cuda_device));
  int rc = upload_domain_table(d);
  if (rc) return rc;
  const bool host_fb = !is_device_ptr(fb);
  float* dfb = fb;
  const size_t img_bytes = sizeof(float) * 4 * (size_t)cam->image_w * cam->image_h;
  const size_t row_bytes = sizeof(float) * 4 * (size_t)cam->image_w;
  const int sb = (int)(d->fb_calls & 1);
  if (host_fb) {
    // Host frame buffer: render into a zeroed device image, then copy the tile's rows back.  (The
    // reference clears the image each frame and tiles are disjoint, src/render/spray_renderer.inl:363.)
    // Two staging images alternate and the copy runs on its own stream: tile k's D2H overlaps tile k+1.
    ++d->fb_calls;
    if (d->fb_stage[sb].bytes < img_bytes + 4096) SG_CUDA(cudaEventSynchronize(d->ev_copy[sb]));  // about to reallocate
    rc = ensure_scratch(d->fb_stage[sb], img_bytes + 4096);
    if (rc) return rc;
    dfb = (float*)d->fb_stage[sb].ptr;
    SG_CUDA(cudaStreamWaitEvent(d->stream, d->ev_copy[sb], 0));   // the copy that last read this staging image
    SG_CUDA(cudaMemsetAsync((char*)dfb + row_bytes * tile.y, 0, row_bytes * tile.h, d->stream));
  }
  rc = launch_render_tile(d, cam, shade, tile, dfb, counted);
  if (rc) return rc;
  if (host_fb && tile.w > 0 && tile.h > 0) {
    const size_t off = (size_t)tile.y * cam->image_w + tile.x;
    SG_CUDA(cudaEventRecord(d->ev_render[sb], d->stream));
    SG_CUDA(cudaStreamWaitEvent(d->copy_stream, d->ev_render[sb], 0));
    SG_CUDA(cudaMemcpy2DAsync(fb + 4 * off, row_bytes, dfb + 4 * off, row_bytes, sizeof(float) * 4 * tile.w, tile.h,
                              cudaMemcpyDeviceToHost, d->copy_stream));
    SG_CUDA(cudaEventRecord(d->ev_copy[sb], d->copy_stream));
  }
  if (stats) {
    unsigned long long* h = (unsigned long long*)d->pinned;
    SG_CUDA(cudaMemcpyAsync(h, d->d_counters, sizeof(unsigned long long) * 8, cudaMemcpyDeviceToHost, d->stream));
    SG_CUDA(cudaStreamSynchronize(d->stream));
    stats->eye_rays = h[1]; stats->radiance_rays = h[2]; stats->shadow_rays = h[3]; stats->domain_tests = h[4];
    stats->nodes_visited = h[5]; stats->tris_tested = h[6]; stats->hits = h[7];
  }
  if (host_fb && !async) SG_CUDA(cudaStreamSynchronize(d->copy_stream));
  return SPRAYGPU_OK;
}

int spraygpu_trace_tile(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade, spraygpu_tile tile,
                        float* fb, spraygpu_stats* stats) {
  return trace_tile_impl(dev, cam, shade, tile, fb, stats, false, false);
}
int spraygpu_trace_tile_async(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                              spraygpu_tile tile, float* fb) {
  return trace_tile_impl(dev, cam, shade, tile, fb, nullptr, false, true);
}
int spraygpu_trace_tile_counted(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                                spraygpu_tile tile, float* fb, spraygpu_stats* stats) {
  return trace_tile_impl(dev, cam, shade, tile, fb, stats, true, false);
}

static int trace_frame_impl(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                            const spraygpu_tile* tiles, const spraygpu_tile* parents, int ntiles, float* fb, spraygpu_stats* stats) {
  SG_CHECK(dev && cam && shade && fb && (tiles || ntiles == 0) && ntiles >= 0, SPRAYGPU_ERR_ARG, "trace_frame: bad argument");
  Device* d = DEV(dev);
  SG_CHECK(!parents || is_device_ptr(fb), SPRAYGPU_ERR_ARG, "trace_frame_rects: fb must be a device pointer");
  if (!is_device_ptr(fb)) {
    // host frame buffer: per-tile launches so that tile k's D2H overlaps tile k+1's kernel
    spraygpu_stats tot = {0, 0, 0, 0, 0, 0, 0}, st;
    for (int t = 0; t < ntiles; ++t) {
      const int rc = trace_tile_impl(dev, cam, shade, tiles[t], fb, stats ? &st : nullptr, false, true);
      if (rc) return rc;
      if (stats) {
        tot.eye_rays += st.eye_rays; tot.radiance_rays += st.radiance_rays; tot.shadow_rays += st.shadow_rays;
        tot.domain_tests += st.domain_tests; tot.hits += st.hits;
      }
    }
    SG_CUDA(cudaStreamSynchronize(d->copy_stream));
    if (stats) *stats = tot;
    return SPRAYGPU_OK;
  }
  SG_CHECK(shade->pixel_samples >= 1 && shade->bounces >= 1, SPRAYGPU_ERR_ARG, "trace_frame: pixel_samples and bounces must be >= 1");
  SG_CHECK(shade->nlights >= 0 && shade->nlights <= SPRAYGPU_MAX_LIGHTS, SPRAYGPU_ERR_LIMIT, "trace_frame: %d lights (limit %d)",
           shade->nlights, SPRAYGPU_MAX_LIGHTS);
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_AO || shade->nlights > 0, SPRAYGPU_ERR_ARG,
           "trace_frame: the PT shader needs at least one light (ooc_tracer.inl:68)");
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_PT || shade->ao_samples > 0, SPRAYGPU_ERR_ARG, "trace_frame: ao_samples must be > 0");
  for (int t = 0; t < ntiles; ++t)
    SG_CHECK(tiles[t].x >= 0 && tiles[t].y >= 0 && tiles[t].w >= 0 && tiles[t].h >= 0 && tiles[t].x + tiles[t].w <= cam->image_w &&
                 tiles[t].y + tiles[t].h <= cam->image_h,
             SPRAYGPU_ERR_ARG, "trace_frame: tile %d (%d,%d,%d,%d) outside the %dx%d image", t, tiles[t].x, tiles[t].y, tiles[t].w,
             tiles[t].h, cam->image_w, cam->image_h);
  SG_CUDA(cudaSetDevice(d->cuda_device));
  int rc = upload_domain_table(d);
  if (rc) return rc;
  const int per = 32;   // SPRAY_MAX_FRAME_TILES
  for (int t0 = 0; t0 < ntiles || t0 == 0; t0 += per) {
    const int nt = ntiles - t0 < per ? ntiles - t0 : per;
    rc = launch_render_tiles(d, cam, shade, tiles + t0, parents ? parents + t0 : nullptr, nt, fb, false, t0 == 0);
    if (rc) return rc;
    if (ntiles == 0) break;
  }
  if (stats) {
    unsigned long long* h = (unsigned long long*)d->pinned;
    SG_CUDA(cudaMemcpyAsync(h, d->d_counters, sizeof(unsigned long long) * 8, cudaMemcpyDeviceToHost, d->stream));
    SG_CUDA(cudaStreamSynchronize(d->stream));
    stats->eye_rays = h[1]; stats->radiance_rays = h[2]; stats->shadow_rays = h[3]; stats->domain_tests = h[4];
    stats->nodes_visited = h[5]; stats->tris_tested = h[6]; stats->hits = h[7];
  }
  return SPRAYGPU_OK;
}

int spraygpu_trace_frame(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                         const spraygpu_tile* tiles, int ntiles, float* fb, spraygpu_stats* stats) {
  return trace_frame_impl(dev, cam, shade, tiles, nullptr, ntiles, fb, stats);
}
int spraygpu_trace_frame_rects(spraygpu_device dev, const spraygpu_camera* cam, const spraygpu_shade* shade,
                               const spraygpu_tile* rects, const spraygpu_tile* blocking_tiles, int n, float* fb,
                               spraygpu_stats* stats) {
  SG_CHECK(blocking_tiles || n == 0, SPRAYGPU_ERR_ARG, "trace_frame_rects: blocking_tiles is NULL");
  return trace_frame_impl(dev, cam, shade, rects, blocking_tiles, n, fb, stats);
}

}  // extern "C"
