// comm.cu — NCCL plumbing of libspraygpu (multi-GPU section of include/spray_gpu.h).
// NCCL is resolved at RUN time (dlopen of libnccl.so.2; an already-loaded copy, e.g. PyTorch's, is
// reused), so the library loads — and the single-GPU path works — on hosts without NCCL.
#include <dlfcn.h>
#include <nccl.h>   // types and enums only; no link-time dependency

#include <algorithm>
#include <mutex>

#include "internal.cuh"

namespace spraygpu {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  const char* (*GetErrorString)(ncclResult_t);
  bool ok = false;
};

static NcclApi g_nccl;
static std::once_flag g_nccl_once;

static void load_nccl() {
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);   // reuse the process's copy if there is one
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
  if (!h) return;
  bool ok = true;
  auto sym = [&](const char* name) -> void* {
    void* p = dlsym(h, name);
    if (!p) ok = false;
    return p;
  };
  g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))sym("ncclGetUniqueId");
  g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))sym("ncclCommInitRank");
  g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))sym("ncclCommDestroy");
  g_nccl.AllReduce = (decltype(g_nccl.AllReduce))sym("ncclAllReduce");
  g_nccl.Reduce = (decltype(g_nccl.Reduce))sym("ncclReduce");
  g_nccl.AllGather = (decltype(g_nccl.AllGather))sym("ncclAllGather");
  g_nccl.Send = (decltype(g_nccl.Send))sym("ncclSend");
  g_nccl.Recv = (decltype(g_nccl.Recv))sym("ncclRecv");
  g_nccl.GroupStart = (decltype(g_nccl.GroupStart))sym("ncclGroupStart");
  g_nccl.GroupEnd = (decltype(g_nccl.GroupEnd))sym("ncclGroupEnd");
  g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))sym("ncclGetErrorString");
  g_nccl.ok = ok;
}

static const NcclApi* nccl() {
  std::call_once(g_nccl_once, load_nccl);
  return g_nccl.ok ? &g_nccl : nullptr;
}

#define SG_NCCL(call)                                                                                  \
  do {                                                                                                 \
    ncclResult_t r__ = (call);                                                                         \
    if (r__ != ncclSuccess) {                                                                          \
      set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r__));        \
      return SPRAYGPU_ERR_CUDA;                                                                        \
    }                                                                                                  \
  } while (0)

// inside an open ncclGroupStart(): close the group before reporting the error, or the communicator's next call hangs
#define SG_NCCL_IN_GROUP(call)                                                                         \
  do {                                                                                                 \
    ncclResult_t r__ = (call);                                                                         \
    if (r__ != ncclSuccess) {                                                                          \
      g_nccl.GroupEnd();                                                                               \
      set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r__));        \
      return SPRAYGPU_ERR_CUDA;                                                                        \
    }                                                                                                  \
  } while (0)

struct Comm {
  Device* dev = nullptr;
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0;
  // state of the peer-memory in-situ path (insitu2.cu), released by spraygpu_comm_destroy
  // and of the ray-forwarding path (insitu3.cu): slot 0 / slot 1
  void* p2p[2] = {nullptr, nullptr};
  void (*p2p_free[2])(Comm*, void*) = {nullptr, nullptr};
};
void** comm_p2p_slot(Comm* c, void (***free_fn)(Comm*, void*), int which) { *free_fn = &c->p2p_free[which]; return &c->p2p[which]; }

// ------------------------------------------------------------------------------------------------
// Peer-mapped device buffers (CUDA IPC over NVLink / NVSwitch P2P): every rank allocates `bytes`, the 64-byte IPC
// handles are all-gathered (setup only, NCCL), and every rank maps every peer's buffer.  After that the frame path
// moves data with plain loads / stores / atomics on the mapped pointers from inside its own kernels — no collective
// call per frame (image.cu: row push; insitu2.cu: key / attribute / occlusion exchange).
// ------------------------------------------------------------------------------------------------
// only_rank >= 0: map that rank's buffer only (the others stay nullptr)
int comm_shared_alloc(Comm* c, size_t bytes, void** local, std::vector<void*>* mapped, int only_rank) {
  SG_CUDA(cudaSetDevice(c->dev->cuda_device));
  *local = nullptr;
  mapped->clear();
  void* p = nullptr;
  char* d_h = nullptr;
  SG_CUDA(cudaMalloc(&p, bytes));
  // every error return below gives back what was taken so far (the caller sees *local == NULL and an empty `mapped`)
  auto fail = [&](int code) {
    for (int r = 0; r < (int)mapped->size(); ++r)
      if (r != c->rank && (*mapped)[(size_t)r]) cudaIpcCloseMemHandle((*mapped)[(size_t)r]);
    mapped->clear();
    if (d_h) cudaFree(d_h);
    cudaFree(p);
    cudaGetLastError();
    return code;
  };
#define SA_CUDA(call)                                                                                       \
  do {                                                                                                      \
    cudaError_t e__ = (call);                                                                               \
    if (e__ != cudaSuccess) {                                                                               \
      set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));                \
      return fail(SPRAYGPU_ERR_CUDA);                                                                       \
    }                                                                                                       \
  } while (0)
  SA_CUDA(cudaMemsetAsync(p, 0, bytes, c->dev->stream));
  if (c->nranks == 1) {
    SA_CUDA(cudaStreamSynchronize(c->dev->stream));
    mapped->assign(1, p);
    *local = p;
    return SPRAYGPU_OK;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t mine;
  SA_CUDA(cudaIpcGetMemHandle(&mine, p));
  SA_CUDA(cudaMalloc((void**)&d_h, 64 * (size_t)c->nranks));
  SA_CUDA(cudaMemcpyAsync(d_h + 64 * (size_t)c->rank, &mine, 64, cudaMemcpyHostToDevice, c->dev->stream));
  {
    ncclResult_t nr = g_nccl.AllGather(d_h + 64 * (size_t)c->rank, d_h, 64, ncclChar, c->comm, c->dev->stream);
    if (nr != ncclSuccess) { set_error("comm_shared_alloc: ncclAllGather failed: %s", g_nccl.GetErrorString(nr)); return fail(SPRAYGPU_ERR_CUDA); }
  }
  std::vector<cudaIpcMemHandle_t> all((size_t)c->nranks);
  SA_CUDA(cudaMemcpyAsync(all.data(), d_h, 64 * (size_t)c->nranks, cudaMemcpyDeviceToHost, c->dev->stream));
  SA_CUDA(cudaStreamSynchronize(c->dev->stream));
  cudaFree(d_h);
  d_h = nullptr;
  mapped->assign((size_t)c->nranks, nullptr);
  for (int r = 0; r < c->nranks; ++r) {
    if (r == c->rank || (only_rank >= 0 && r != only_rank)) continue;
    void* q = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&q, all[(size_t)r], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      set_error("cudaIpcOpenMemHandle(rank %d) failed: %s (peer access over NVLink is required)", r, cudaGetErrorString(e));
      return fail(SPRAYGPU_ERR_CUDA);
    }
    (*mapped)[(size_t)r] = q;
  }
#undef SA_CUDA
  (*mapped)[(size_t)c->rank] = p;
  *local = p;
  return SPRAYGPU_OK;
}
// Collective agreement on a set-up step: returns 1 only if EVERY rank passed ok != 0 (min over ranks), so that all ranks
// take the same branch afterwards — one rank bailing out alone would leave its peers in the next collective for ever.
// Also a rendezvous (stream sync + all-reduce), like comm_host_barrier.  -1: the exchange itself failed.
int comm_all_ok(Comm* c, int ok) {
  if (c->nranks == 1) return ok ? 1 : 0;
  if (cudaSetDevice(c->dev->cuda_device) != cudaSuccess) return -1;
  unsigned long long v = ok ? 1ull : 0ull;
  unsigned long long* d = c->dev->d_counters + 9;        // scratch word no kernel uses
  if (cudaMemcpyAsync(d, &v, sizeof(v), cudaMemcpyHostToDevice, c->dev->stream) != cudaSuccess) return -1;
  if (g_nccl.AllReduce(d, d, 1, ncclUint64, ncclMin, c->comm, c->dev->stream) != ncclSuccess) return -1;
  if (cudaMemcpyAsync(&v, d, sizeof(v), cudaMemcpyDeviceToHost, c->dev->stream) != cudaSuccess) return -1;
  if (cudaStreamSynchronize(c->dev->stream) != cudaSuccess) return -1;
  return v ? 1 : 0;
}
// collective: every rank must have stopped using the peers' mappings (the caller synchronises ranks first)
void comm_shared_free(Comm* c, void* local, std::vector<void*>* mapped) {
  cudaSetDevice(c->dev->cuda_device);
  for (int r = 0; r < (int)mapped->size(); ++r)
    if (r != c->rank && (*mapped)[(size_t)r]) cudaIpcCloseMemHandle((*mapped)[(size_t)r]);
  mapped->clear();
  if (local) cudaFree(local);
}
// host-visible rendezvous of all ranks (setup / teardown only): a 1-int all-reduce + stream sync
int comm_host_barrier(Comm* c) {
  if (c->nranks == 1) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(c->dev->cuda_device));
  SG_NCCL(g_nccl.AllReduce(c->dev->d_counters, c->dev->d_counters, 1, ncclUint64, ncclSum, c->comm, c->dev->stream));
  SG_CUDA(cudaStreamSynchronize(c->dev->stream));
  return SPRAYGPU_OK;
}

// collectives used by insitu.cu (all on dev->stream, in place)
int comm_allreduce_u64_min(Comm* c, uint64_t* p, size_t n) {
  if (c->nranks == 1 || !n) return SPRAYGPU_OK;
  SG_NCCL(g_nccl.AllReduce(p, p, n, ncclUint64, ncclMin, c->comm, c->dev->stream));
  return SPRAYGPU_OK;
}
int comm_allreduce_u8_max(Comm* c, uint8_t* p, size_t n) {
  if (c->nranks == 1 || !n) return SPRAYGPU_OK;
  SG_NCCL(g_nccl.AllReduce(p, p, n, ncclUint8, ncclMax, c->comm, c->dev->stream));
  return SPRAYGPU_OK;
}
// Buffers with DISJOINT support across ranks (exactly one rank writes a non-zero word, the rest hold 0):
// an integer sum moves the bit patterns unchanged (a float sum would turn -0.0 into +0.0).
int comm_allreduce_bits_sum(Comm* c, uint32_t* p, size_t nwords) {
  if (c->nranks == 1 || !nwords) return SPRAYGPU_OK;
  SG_NCCL(g_nccl.AllReduce(p, p, nwords, ncclUint32, ncclSum, c->comm, c->dev->stream));
  return SPRAYGPU_OK;
}
int comm_allreduce_u64_sum(Comm* c, unsigned long long* p, size_t n) {
  if (c->nranks == 1 || !n) return SPRAYGPU_OK;
  SG_NCCL(g_nccl.AllReduce(p, p, n, ncclUint64, ncclSum, c->comm, c->dev->stream));
  return SPRAYGPU_OK;
}
int comm_rank(const Comm* c) { return c->rank; }
int comm_size(const Comm* c) { return c->nranks; }
Device* comm_device(const Comm* c) { return c->dev; }

}  // namespace spraygpu

using namespace spraygpu;

extern "C" {

int spraygpu_comm_unique_id(char id[SPRAYGPU_UNIQUE_ID_BYTES]) {
  SG_CHECK(id, SPRAYGPU_ERR_ARG, "comm_unique_id: id is NULL");
  SG_CHECK(nccl(), SPRAYGPU_ERR_STATE, "NCCL (libnccl.so.2) is not available on this host");
  static_assert(sizeof(ncclUniqueId) == SPRAYGPU_UNIQUE_ID_BYTES, "unique id size");
  ncclUniqueId u;
  SG_NCCL(g_nccl.GetUniqueId(&u));
  memcpy(id, &u, sizeof(u));
  return SPRAYGPU_OK;
}

int spraygpu_comm_create(spraygpu_device dev, int nranks, int rank, const char id[SPRAYGPU_UNIQUE_ID_BYTES],
                         spraygpu_comm* out) {
  SG_CHECK(dev && out && id, SPRAYGPU_ERR_ARG, "comm_create: NULL argument");
  SG_CHECK(nranks >= 1 && rank >= 0 && rank < nranks, SPRAYGPU_ERR_ARG, "comm_create: bad rank %d of %d", rank, nranks);
  SG_CHECK(nccl(), SPRAYGPU_ERR_STATE, "NCCL (libnccl.so.2) is not available on this host");
  Device* d = (Device*)dev;
  SG_CUDA(cudaSetDevice(d->cuda_device));
  Comm* c = new Comm();
  c->dev = d; c->nranks = nranks; c->rank = rank;
  ncclUniqueId u;
  memcpy(&u, id, sizeof(u));
  ncclResult_t r = g_nccl.CommInitRank(&c->comm, nranks, u, rank);
  if (r != ncclSuccess) {
    set_error("ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
    delete c;
    return SPRAYGPU_ERR_CUDA;
  }
  *out = (spraygpu_comm)c;
  return SPRAYGPU_OK;
}

int spraygpu_comm_destroy(spraygpu_comm comm) {
  if (!comm) return SPRAYGPU_OK;
  Comm* c = (Comm*)comm;
  cudaSetDevice(c->dev->cuda_device);
  cudaStreamSynchronize(c->dev->stream);
  for (int k = 0; k < 2; ++k)
    if (c->p2p[k] && c->p2p_free[k]) { c->p2p_free[k](c, c->p2p[k]); c->p2p[k] = nullptr; }   // collective (host barrier inside)
  if (c->comm && nccl()) g_nccl.CommDestroy(c->comm);
  delete c;
  return SPRAYGPU_OK;
}

// InsituPartition::partition, src/render/data_partition.h:59-137 (GROUP_CLOSE_DOMAINS branch) and :139-155
int spraygpu_partition_domains(int nd, const float* wb, const float sb[6], int nranks, int round_robin, int32_t* d2r,
                               uint32_t* codes_out) {
  SG_CHECK(nd > 0 && wb && sb && d2r && nranks > 0, SPRAYGPU_ERR_ARG, "partition_domains: bad argument");
  // data_partition.h:71-77: scale = 1 / diagonal, offset = 0 - lo * scale (per component, FP32)
  float sf[3], off[3];
  for (int a = 0; a < 3; ++a) {
    const float diag = sb[3 + a] - sb[a];
    sf[a] = 1.0f / diag;
    off[a] = 0.0f - sb[a] * sf[a];
  }
  std::vector<std::pair<uint32_t, int>> order(nd);
  for (int i = 0; i < nd; ++i) {
    float p[3];
    for (int a = 0; a < 3; ++a) {
      const float c = (wb[6 * i + a] + wb[6 * i + 3 + a]) * 0.5f;   // Aabb::getCenter
      p[a] = sf[a] * c + off[a];
    }
    order[i] = {morton30(p[0], p[1], p[2]), i};
    if (codes_out) codes_out[i] = order[i].first;
  }
  // the reference's std::sort is unstable on equal codes; ties are broken by domain id here
  std::sort(order.begin(), order.end());
  if (round_robin) {
    for (int k = 0; k < nd; ++k) d2r[order[k].second] = k % nranks;
    return SPRAYGPU_OK;
  }
  const int shares = std::max(nd / nranks, 1);
  int rank = 0, s = 0;
  for (int k = 0; k < nd; ++k) {
    d2r[order[k].second] = rank;
    if (++s == shares) {
      s = 0;
      if (++rank == nranks) rank = 0;
    }
  }
  return SPRAYGPU_OK;
}

int spraygpu_image_reduce(spraygpu_device dev, spraygpu_comm comm, float* fb, size_t npixels, int root) {
  SG_CHECK(dev && comm && fb, SPRAYGPU_ERR_ARG, "image_reduce: NULL argument");
  Comm* c = (Comm*)comm;
  SG_CHECK(root >= 0 && root < c->nranks, SPRAYGPU_ERR_ARG, "image_reduce: bad root %d", root);
  if (c->nranks == 1 || !npixels) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(c->dev->cuda_device));
  SG_NCCL(g_nccl.Reduce(fb, fb, npixels * 4, ncclFloat32, ncclSum, root, c->comm, c->dev->stream));
  return SPRAYGPU_OK;
}

int spraygpu_image_gather_rows(spraygpu_device dev, spraygpu_comm comm, float* fb, int image_w, const int32_t* row_begin,
                               int root) {
  SG_CHECK(dev && comm && fb && row_begin && image_w > 0, SPRAYGPU_ERR_ARG, "image_gather_rows: bad argument");
  Comm* c = (Comm*)comm;
  SG_CHECK(root >= 0 && root < c->nranks, SPRAYGPU_ERR_ARG, "image_gather_rows: bad root %d", root);
  if (c->nranks == 1) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(c->dev->cuda_device));
  const size_t row_floats = (size_t)image_w * 4;
  SG_NCCL(g_nccl.GroupStart());
  if (c->rank == root) {
    for (int r = 0; r < c->nranks; ++r) {
      if (r == root) continue;
      const size_t n = (size_t)(row_begin[r + 1] - row_begin[r]) * row_floats;
      if (n) SG_NCCL_IN_GROUP(g_nccl.Recv(fb + (size_t)row_begin[r] * row_floats, n, ncclFloat32, r, c->comm, c->dev->stream));
    }
  } else {
    const size_t n = (size_t)(row_begin[c->rank + 1] - row_begin[c->rank]) * row_floats;
    if (n) SG_NCCL_IN_GROUP(g_nccl.Send(fb + (size_t)row_begin[c->rank] * row_floats, n, ncclFloat32, root, c->comm, c->dev->stream));
  }
  SG_NCCL(g_nccl.GroupEnd());
  return SPRAYGPU_OK;
}

int spraygpu_image_gather_blocks(spraygpu_device dev, spraygpu_comm comm, float* fb, int image_w, int nblocks,
                                 const int32_t* block_row_begin, const int32_t* block_owner, int root) {
  SG_CHECK(dev && comm && fb && block_row_begin && block_owner && image_w > 0 && nblocks >= 0, SPRAYGPU_ERR_ARG,
           "image_gather_blocks: bad argument");
  Comm* c = (Comm*)comm;
  SG_CHECK(root >= 0 && root < c->nranks, SPRAYGPU_ERR_ARG, "image_gather_blocks: bad root %d", root);
  if (c->nranks == 1) return SPRAYGPU_OK;
  SG_CUDA(cudaSetDevice(c->dev->cuda_device));
  const size_t row_floats = (size_t)image_w * 4;
  for (int b = 0; b < nblocks; ++b)   // validated BEFORE the group opens: no early return with an open group
    SG_CHECK(block_owner[b] >= 0 && block_owner[b] < c->nranks, SPRAYGPU_ERR_ARG, "image_gather_blocks: block %d owner %d", b, block_owner[b]);
  SG_NCCL(g_nccl.GroupStart());
  for (int b = 0; b < nblocks; ++b) {
    const int owner = block_owner[b];
    const size_t n = (size_t)(block_row_begin[b + 1] - block_row_begin[b]) * row_floats;
    if (!n || owner == root) continue;
    float* p = fb + (size_t)block_row_begin[b] * row_floats;
    if (c->rank == root) SG_NCCL_IN_GROUP(g_nccl.Recv(p, n, ncclFloat32, owner, c->comm, c->dev->stream));
    else if (c->rank == owner) SG_NCCL_IN_GROUP(g_nccl.Send(p, n, ncclFloat32, root, c->comm, c->dev->stream));
  }
  SG_NCCL(g_nccl.GroupEnd());
  return SPRAYGPU_OK;
}

int spraygpu_allreduce_keys_min(spraygpu_device dev, spraygpu_comm comm, spraygpu_key* keys, size_t n) {
  SG_CHECK(dev && comm && (keys || !n), SPRAYGPU_ERR_ARG, "allreduce_keys_min: NULL argument");
  SG_CUDA(cudaSetDevice(((Comm*)comm)->dev->cuda_device));
  return comm_allreduce_u64_min((Comm*)comm, keys, n);
}
int spraygpu_allreduce_occ_max(spraygpu_device dev, spraygpu_comm comm, uint8_t* occ, size_t n) {
  SG_CHECK(dev && comm && (occ || !n), SPRAYGPU_ERR_ARG, "allreduce_occ_max: NULL argument");
  SG_CUDA(cudaSetDevice(((Comm*)comm)->dev->cuda_device));
  return comm_allreduce_u8_max((Comm*)comm, occ, n);
}

}  // extern "C"
