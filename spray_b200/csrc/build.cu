// build.cu — per-domain acceleration-structure build on the GPU (= what rtcCommit does for the
// reference, src/render/trimesh_buffer.cc:189-225), plus the vertex transform (:141-157) and the
// per-vertex normals (:267-326).
//
// Pipeline (all on dev->stream, no host round trips except one 8-byte counter read per BVH8 level):
//   transform -> per-vertex face lists (counting scatter) sorted per vertex -> deterministic per-vertex normal sums
//   -> triangle boxes + centroid bounds -> 30-bit Morton keys -> LSD radix sort
//   -> Karras LBVH -> bottom-up refit -> level-synchronous collapse into 80-byte 8-wide nodes
//   -> triangle records (48 B, leaf order) + attribute records (48 B, primID order)
#include <cooperative_groups.h>
#include <stdlib.h>

#include <algorithm>

#include "internal.cuh"

namespace cg = cooperative_groups;

namespace spraygpu {

// ------------------------------------------------------------------------------------------------
// LSD radix sort, 8 bits per pass, stable, key+value (uint32).  Tile = 256 threads x 16 keys.
// ------------------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;                      // keys per thread of a full-size tile
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
// A tile is worked off in rounds of 256 keys with block-wide barriers between them, so a SMALL input sorted with
// full-size tiles is pure latency (ncu r04f: 17 us per scatter pass for the 5 480-triangle wavelet domain, two CTAs on
// the whole GPU).  Small inputs get fewer keys per thread, i.e. more CTAs with fewer rounds each.
static inline int rs_items_for(int n, int sm_count) {
  const int want = (n + RS_THREADS * 2 * sm_count - 1) / (RS_THREADS * 2 * sm_count);
  return want < 1 ? 1 : (want > RS_ITEMS ? RS_ITEMS : want);
}

__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint32_t* __restrict__ keys, int n, int shift,
                                                             uint32_t* __restrict__ hist, int nblocks, int items) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int base = blockIdx.x * RS_THREADS * items;
  for (int r = 0; r < items; ++r) {
    int i = base + r * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&h[(keys[i] >> shift) & 0xffu], 1u);
  }
  __syncthreads();
  hist[threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// Exclusive scan of the `total` (= 256 * nblocks, a multiple of 4) digit counters in two small kernels: every block
// scans one chunk of 4096 counters (coalesced 16-byte loads, shuffle scan) and leaves the chunk total in sums[];
// one block then scans the <= 1024 chunk totals.  The scatter kernel adds sums[chunk] when it reads its bases.
// (The first version scanned everything with ONE block and strided per-thread loops: ncu r01u showed it at 53 % of
// the whole commit for a 1 Mi-triangle domain.)
constexpr int SC_CHUNK = 4096;
__global__ void __launch_bounds__(1024) rs_scan_chunks_kernel(uint32_t* __restrict__ hist, int total, uint32_t* __restrict__ sums) {
  __shared__ uint32_t wsum[32];
  const int i0 = blockIdx.x * SC_CHUNK + threadIdx.x * 4;
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  if (i0 < total) v = *reinterpret_cast<const uint4*>(hist + i0);
  const uint32_t t = v.x + v.y + v.z + v.w;
  uint32_t inc = t;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += o;
  }
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = wsum[lane], wi = w;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, wi, off);
      if (lane >= off) wi += o;
    }
    wsum[lane] = wi - w;   // exclusive warp offsets
    if (lane == 31) sums[blockIdx.x] = wi;
  }
  __syncthreads();
  uint32_t run = wsum[warp] + inc - t;
  if (i0 < total) {
    uint4 o;
    o.x = run; run += v.x;
    o.y = run; run += v.y;
    o.z = run; run += v.z;
    o.w = run;
    *reinterpret_cast<uint4*>(hist + i0) = o;
  }
}
__global__ void __launch_bounds__(1024) rs_scan_sums_kernel(uint32_t* __restrict__ sums, int nchunks) {
  __shared__ uint32_t wsum[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t t = (int)threadIdx.x < nchunks ? sums[threadIdx.x] : 0u;
  uint32_t inc = t;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += o;
  }
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = wsum[lane], wi = w;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, wi, off);
      if (lane >= off) wi += o;
    }
    wsum[lane] = wi - w;
  }
  __syncthreads();
  if ((int)threadIdx.x < nchunks) sums[threadIdx.x] = wsum[warp] + inc - t;
}

__global__ void __launch_bounds__(RS_THREADS)
    rs_scatter_kernel(const uint32_t* __restrict__ kin, const uint32_t* __restrict__ vin, uint32_t* __restrict__ kout,
                      uint32_t* __restrict__ vout, int n, int shift, const uint32_t* __restrict__ hist, int nblocks,
                      const uint32_t* __restrict__ sums, int items) {
  __shared__ uint32_t digit_base[256];
  __shared__ uint32_t warp_cnt[RS_THREADS / 32][256];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  {
    const int idx = tid * nblocks + blockIdx.x;
    digit_base[tid] = hist[idx] + sums[idx / SC_CHUNK];
  }
  const int base = blockIdx.x * RS_THREADS * items;
  for (int r = 0; r < items; ++r) {
#pragma unroll
    for (int w = 0; w < RS_THREADS / 32; ++w) warp_cnt[w][tid] = 0;
    __syncthreads();
    const int i = base + r * RS_THREADS + tid;
    const bool valid = i < n;
    uint32_t key = 0, val = 0, d = 0x100u;  // invalid lanes share a private digit
    if (valid) { key = kin[i]; val = vin[i]; d = (key >> shift) & 0xffu; }
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    if (valid && rank == 0) warp_cnt[warp][d] = __popc(peers);
    __syncthreads();
    uint32_t tot = 0;
#pragma unroll
    for (int w = 0; w < RS_THREADS / 32; ++w) {
      uint32_t c = warp_cnt[w][tid];
      warp_cnt[w][tid] = tot;
      tot += c;
    }
    __syncthreads();
    if (valid) {
      uint32_t pos = digit_base[d] + warp_cnt[warp][d] + rank;
      kout[pos] = key;
      vout[pos] = val;
    }
    __syncthreads();
    digit_base[tid] += tot;
  }
}

// sorts (keys, vals) of length n on `bits` low key bits; result ends in (*k, *v) (pointers may swap)
static int radix_sort(Device* dev, cudaStream_t st, uint32_t** k, uint32_t** v, uint32_t** k2, uint32_t** v2, uint32_t* hist,
                      int n, int bits) {
  if (n <= 1) return SPRAYGPU_OK;
  const int items = rs_items_for(n, dev->sm_count);
  const int nblocks = (n + RS_THREADS * items - 1) / (RS_THREADS * items);
  const int passes = (bits + 7) / 8;
  const int total = 256 * nblocks, nchunks = (total + SC_CHUNK - 1) / SC_CHUNK;
  SG_CHECK(nchunks <= 1024, SPRAYGPU_ERR_LIMIT, "radix sort: %d keys exceed the two-level scan (max %d)", n, 1024 * SC_CHUNK / 256 * RS_TILE);
  uint32_t* sums = hist + (size_t)256 * nblocks;   // 1024 chunk totals live behind the counters
  for (int p = 0; p < passes; ++p) {
    const int shift = 8 * p;
    rs_hist_kernel<<<nblocks, RS_THREADS, 0, st>>>(*k, n, shift, hist, nblocks, items);
    rs_scan_chunks_kernel<<<nchunks, 1024, 0, st>>>(hist, total, sums);
    rs_scan_sums_kernel<<<1, 1024, 0, st>>>(sums, nchunks);
    rs_scatter_kernel<<<nblocks, RS_THREADS, 0, st>>>(*k, *v, *k2, *v2, n, shift, hist, nblocks, sums, items);
    dev->launches += 4;
    std::swap(*k, *k2);
    std::swap(*v, *v2);
  }
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

// ------------------------------------------------------------------------------------------------
// geometry kernels
// ------------------------------------------------------------------------------------------------
struct Mat4 { float m[16]; };

// TriMeshBuffer::load, trimesh_buffer.cc:141-157: v = x * vec4(v,1) with glm's mat4*vec4 order
// (deps/glm/glm/detail/type_mat4x4.inl:536-): (m0*x + m1*y) + (m2*z + m3*1), no contraction.
__global__ void transform_kernel(const float* __restrict__ vin, float* __restrict__ vout, size_t nv, Mat4 M) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= nv) return;
  float x = vin[3 * i], y = vin[3 * i + 1], z = vin[3 * i + 2];
  const float* m = M.m;
  vout[3 * i + 0] = add_rn(add_rn(mul_rn(m[0], x), mul_rn(m[4], y)), add_rn(mul_rn(m[8], z), mul_rn(m[12], 1.0f)));
  vout[3 * i + 1] = add_rn(add_rn(mul_rn(m[1], x), mul_rn(m[5], y)), add_rn(mul_rn(m[9], z), mul_rn(m[13], 1.0f)));
  vout[3 * i + 2] = add_rn(add_rn(mul_rn(m[2], x), mul_rn(m[6], y)), add_rn(mul_rn(m[10], z), mul_rn(m[14], 1.0f)));
}

// ------------------------------------------------------------------------------------------------
// Per-vertex normals WITHOUT the (vertex, face) pair sort (round 2).  The reference adds the face normals to a vertex
// in increasing face order (trimesh_buffer.cc:267-326) and the float sum depends on that order, so each vertex needs
// its incident faces SORTED — but only within the vertex.  Round 1 radix-sorted all 3 n pairs by vertex (3 passes over
// 3 Mi pairs for a 1 Mi-triangle domain: ~280 us of a 1.06 ms commit).  Here: count the valence, hand every vertex a
// segment (block scan + one atomic per 1024 vertices: segments need no particular order), scatter the face ids with
// per-vertex atomics (arbitrary order), and let the summing thread sort its own short list (valence <= 32: insertion
// sort in local memory; heavier vertices — the apex of a fan — go to a list worked off by one CTA each with an
// all-ascending bitonic network in place).  Same faces, same order, bit-identical sums.
// ------------------------------------------------------------------------------------------------
#define NRM_LIGHT_MAX 32
__global__ void vcount_kernel(const uint32_t* __restrict__ faces, int npairs, uint32_t* __restrict__ vcnt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < npairs) atomicAdd(&vcnt[faces[i]], 1u);
}
__global__ void __launch_bounds__(1024) valloc_kernel(const uint32_t* __restrict__ vcnt, size_t nv, uint32_t* __restrict__ vstart,
                                                      unsigned int* __restrict__ cursor) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t base_s;
  const size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const uint32_t c = v < nv ? vcnt[v] : 0u;
  uint32_t inc = c;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += o;
  }
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const uint32_t w = wsum[lane];
    uint32_t wi = w;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, wi, off);
      if (lane >= off) wi += o;
    }
    wsum[lane] = wi - w;
    if (lane == 31) base_s = wi ? atomicAdd(cursor, wi) : 0u;
  }
  __syncthreads();
  if (v < nv) vstart[v] = base_s + wsum[warp] + inc - c;
}
__global__ void vfill_kernel(const uint32_t* __restrict__ faces, int npairs, const uint32_t* __restrict__ vstart,
                             uint32_t* __restrict__ vfill, uint32_t* __restrict__ vlist) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npairs) return;
  const uint32_t v = faces[i];
  vlist[vstart[v] + atomicAdd(&vfill[v], 1u)] = (uint32_t)i / 3u;
}
__device__ __forceinline__ void face_normal(const float* __restrict__ verts, const uint32_t* __restrict__ faces, uint32_t f, float* fx,
                                            float* fy, float* fz) {
  const uint32_t i0 = faces[3 * (size_t)f], i1 = faces[3 * (size_t)f + 1], i2 = faces[3 * (size_t)f + 2];
  const vec3 v0 = mk3(verts[3 * (size_t)i0], verts[3 * (size_t)i0 + 1], verts[3 * (size_t)i0 + 2]);
  const vec3 v1 = mk3(verts[3 * (size_t)i1], verts[3 * (size_t)i1 + 1], verts[3 * (size_t)i1 + 2]);
  const vec3 v2 = mk3(verts[3 * (size_t)i2], verts[3 * (size_t)i2 + 1], verts[3 * (size_t)i2 + 2]);
  const vec3 u = gsub(v1, v0), w = gsub(v2, v0);
  *fx = sub_rn(mul_rn(u.y, w.z), mul_rn(u.z, w.y));
  *fy = sub_rn(mul_rn(u.z, w.x), mul_rn(u.x, w.z));
  *fz = sub_rn(mul_rn(u.x, w.y), mul_rn(u.y, w.x));
}
__global__ void normals_light_kernel(const float* __restrict__ verts, const uint32_t* __restrict__ faces,
                                     const uint32_t* __restrict__ vlist, const uint32_t* __restrict__ vstart,
                                     const uint32_t* __restrict__ vcnt, float* __restrict__ normals, size_t nv,
                                     uint32_t* __restrict__ heavy, unsigned int* __restrict__ heavy_n) {
  const size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= nv) return;
  const uint32_t d = vcnt[v], b = vstart[v];
  if (d > NRM_LIGHT_MAX) { heavy[atomicAdd(heavy_n, 1u)] = (uint32_t)v; return; }
  uint32_t fl[NRM_LIGHT_MAX];
  for (uint32_t k = 0; k < d; ++k) {           // insertion sort while loading
    const uint32_t f = vlist[b + k];
    uint32_t j = k;
    while (j > 0 && fl[j - 1] > f) { fl[j] = fl[j - 1]; --j; }
    fl[j] = f;
  }
  float nx = 0.0f, ny = 0.0f, nz = 0.0f;
  for (uint32_t k = 0; k < d; ++k) {
    float fx, fy, fz;
    face_normal(verts, faces, fl[k], &fx, &fy, &fz);
    nx = add_rn(nx, fx); ny = add_rn(ny, fy); nz = add_rn(nz, fz);
  }
  normals[3 * v] = nx; normals[3 * v + 1] = ny; normals[3 * v + 2] = nz;
}
// one CTA per heavy vertex: sort its face list in place (bitonic network with every comparator ascending, so the virtual
// +inf padding up to the next power of two never moves), then sum in order — face normals in parallel per 256-entry
// chunk, the additions by one thread (the order IS the result)
__global__ void __launch_bounds__(256) normals_heavy_kernel(const float* __restrict__ verts, const uint32_t* __restrict__ faces,
                                                            uint32_t* __restrict__ vlist, const uint32_t* __restrict__ vstart,
                                                            const uint32_t* __restrict__ vcnt, float* __restrict__ normals,
                                                            const uint32_t* __restrict__ heavy, const unsigned int* __restrict__ heavy_n) {
  __shared__ float fn[256][3];
  const unsigned int nh = *heavy_n;
  for (unsigned int h = blockIdx.x; h < nh; h += gridDim.x) {
    const uint32_t v = heavy[h], d = vcnt[v];
    uint32_t* a = vlist + vstart[v];
    uint32_t P = 1;
    while (P < d) P <<= 1;
    for (uint32_t k = 2; k <= P; k <<= 1) {
      for (uint32_t j = k >> 1; j > 0; j >>= 1) {
        for (uint32_t i = threadIdx.x; i < d; i += blockDim.x) {
          const uint32_t l = (j == (k >> 1)) ? (i ^ (k - 1)) : (i ^ j);
          if (l > i && l < d) {
            const uint32_t x = a[i], y = a[l];
            if (x > y) { a[i] = y; a[l] = x; }
          }
        }
        __syncthreads();
      }
    }
    float nx = 0.0f, ny = 0.0f, nz = 0.0f;
    for (uint32_t c0 = 0; c0 < d; c0 += 256) {
      const uint32_t k = c0 + threadIdx.x;
      if (k < d) face_normal(verts, faces, a[k], &fn[threadIdx.x][0], &fn[threadIdx.x][1], &fn[threadIdx.x][2]);
      __syncthreads();
      if (threadIdx.x == 0) {
        const uint32_t m = min(256u, d - c0);
        for (uint32_t q = 0; q < m; ++q) { nx = add_rn(nx, fn[q][0]); ny = add_rn(ny, fn[q][1]); nz = add_rn(nz, fn[q][2]); }
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) { normals[3 * (size_t)v] = nx; normals[3 * (size_t)v + 1] = ny; normals[3 * (size_t)v + 2] = nz; }
    __syncthreads();
  }
}

__device__ __forceinline__ void atomic_min_f(uint32_t* a, float v) { atomicMin(a, float_sortable(v)); }
__device__ __forceinline__ void atomic_max_f(uint32_t* a, float v) { atomicMax(a, float_sortable(v)); }

// per-triangle box; block-reduced bounds of (a) centroids (Morton frame) and (b) geometry
__global__ void __launch_bounds__(256)
    tri_bounds_kernel(const float* __restrict__ verts, const uint32_t* __restrict__ faces, int nf,
                      float* __restrict__ tlo, float* __restrict__ thi, uint32_t* __restrict__ bounds /*12*/) {
  // grid-stride, per-thread running bounds, then warp -> block -> 12 atomics per BLOCK (the first version issued 12
  // contended atomics per warp: 270 us for 1 Mi triangles, ncu r01u)
  float cl[3] = {INFINITY, INFINITY, INFINITY}, ch[3] = {-INFINITY, -INFINITY, -INFINITY};
  float gl[3] = {INFINITY, INFINITY, INFINITY}, gh[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nf; i += gridDim.x * blockDim.x) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const uint32_t vi = faces[3 * (size_t)i + k];
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        float x = verts[3 * (size_t)vi + a];
        lo[a] = fminf(lo[a], x); hi[a] = fmaxf(hi[a], x);
      }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      tlo[3 * (size_t)i + a] = lo[a]; thi[3 * (size_t)i + a] = hi[a];
      const float c = 0.5f * (lo[a] + hi[a]);
      cl[a] = fminf(cl[a], c); ch[a] = fmaxf(ch[a], c);
      gl[a] = fminf(gl[a], lo[a]); gh[a] = fmaxf(gh[a], hi[a]);
    }
  }
  __shared__ float red[8][12];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int off = 16; off; off >>= 1) {
      cl[a] = fminf(cl[a], __shfl_xor_sync(0xffffffffu, cl[a], off));
      ch[a] = fmaxf(ch[a], __shfl_xor_sync(0xffffffffu, ch[a], off));
      gl[a] = fminf(gl[a], __shfl_xor_sync(0xffffffffu, gl[a], off));
      gh[a] = fmaxf(gh[a], __shfl_xor_sync(0xffffffffu, gh[a], off));
    }
    if (lane == 0) { red[warp][a] = cl[a]; red[warp][3 + a] = ch[a]; red[warp][6 + a] = gl[a]; red[warp][9 + a] = gh[a]; }
  }
  __syncthreads();
  if (threadIdx.x < 12) {
    const int k = threadIdx.x;
    const bool is_min = k < 3 || (k >= 6 && k < 9);
    float v = red[0][k];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) v = is_min ? fminf(v, red[w][k]) : fmaxf(v, red[w][k]);
    if (is_min) atomic_min_f(&bounds[k], v); else atomic_max_f(&bounds[k], v);
  }
}

__global__ void morton_kernel(const float* __restrict__ tlo, const float* __restrict__ thi, int nf,
                              const uint32_t* __restrict__ bounds, uint32_t* __restrict__ keys,
                              uint32_t* __restrict__ vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nf) return;
  float p[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float lo = sortable_float(bounds[a]), hi = sortable_float(bounds[3 + a]);
    float c = 0.5f * (tlo[3 * (size_t)i + a] + thi[3 * (size_t)i + a]);
    float ext = hi - lo;
    p[a] = ext > 0.0f ? (c - lo) / ext : 0.0f;
  }
  keys[i] = morton30(p[0], p[1], p[2]);
  vals[i] = (uint32_t)i;
}

__global__ void leaf_boxes_kernel(const uint32_t* __restrict__ sorted, const float* __restrict__ tlo,
                                  const float* __restrict__ thi, int n, float* __restrict__ llo,
                                  float* __restrict__ lhi) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const size_t t = sorted[i];
#pragma unroll
  for (int a = 0; a < 3; ++a) { llo[3 * (size_t)i + a] = tlo[3 * t + a]; lhi[3 * (size_t)i + a] = thi[3 * t + a]; }
}

__global__ void lbvh_kernel(const uint32_t* __restrict__ keys, int n, int* __restrict__ bl, int* __restrict__ br,
                            int* __restrict__ bfirst, int* __restrict__ blast, int* __restrict__ parent) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  int l, r, f, la;
  lbvh_node(keys, n, i, &l, &r, &f, &la);
  bl[i] = l; br[i] = r; bfirst[i] = f; blast[i] = la;
  parent[l < 0 ? (n - 1 + ~l) : l] = i;
  parent[r < 0 ? (n - 1 + ~r) : r] = i;
  if (i == 0) parent[0] = -1;
}

// also leaves, per internal node, the 16-byte record the collapse walks: (left, right, leaf count, half-area bits)
__global__ void refit_kernel(const int* __restrict__ parent, const int* __restrict__ bl, const int* __restrict__ br,
                             const int* __restrict__ bfirst, const int* __restrict__ blast,
                             const float* __restrict__ llo, const float* __restrict__ lhi, float* blo, float* bhi,
                             int4* __restrict__ info, int* flags, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int p = parent[n - 1 + i];
  while (p >= 0) {
    __threadfence();
    if (atomicAdd(&flags[p], 1) == 0) return;  // first arrival: the sibling subtree is not done yet
    __threadfence();
    const int cs[2] = {bl[p], br[p]};
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int c = cs[k];
      const volatile float* slo = c < 0 ? llo + 3 * (size_t)(~c) : blo + 3 * (size_t)c;
      const volatile float* shi = c < 0 ? lhi + 3 * (size_t)(~c) : bhi + 3 * (size_t)c;
#pragma unroll
      for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], slo[a]); hi[a] = fmaxf(hi[a], shi[a]); }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) { blo[3 * (size_t)p + a] = lo[a]; bhi[3 * (size_t)p + a] = hi[a]; }
    info[p] = make_int4(cs[0], cs[1], blast[p] - bfirst[p] + 1, __float_as_int(box_half_area(lo, hi)));
    p = parent[p];
  }
}

__device__ __forceinline__ void write_tri(const float* __restrict__ verts, const uint32_t* __restrict__ faces,
                                          uint32_t prim, float4* __restrict__ out) {
  const uint32_t i0 = faces[3 * (size_t)prim], i1 = faces[3 * (size_t)prim + 1], i2 = faces[3 * (size_t)prim + 2];
  vec3 v0 = mk3(verts[3 * (size_t)i0], verts[3 * (size_t)i0 + 1], verts[3 * (size_t)i0 + 2]);
  vec3 v1 = mk3(verts[3 * (size_t)i1], verts[3 * (size_t)i1 + 1], verts[3 * (size_t)i1 + 2]);
  vec3 v2 = mk3(verts[3 * (size_t)i2], verts[3 * (size_t)i2 + 1], verts[3 * (size_t)i2 + 2]);
  vec3 e1 = gsub(v0, v1), e2 = gsub(v2, v0);
  out[0] = make_float4(v0.x, v0.y, v0.z, __uint_as_float(prim));
  out[1] = make_float4(e1.x, e1.y, e1.z, 0.0f);
  out[2] = make_float4(e2.x, e2.y, e2.z, 0.0f);
}

// ------------------------------------------------------------------------------------------------
// Collapse: binary LBVH -> 80-byte 8-wide nodes, ONE persistent cooperative launch.
//
// Round 1 ran one kernel per BVH8 level with one heavy thread per node (collapse_select + encode_node8: ~1 KB of
// local arrays) and a counter read-back + cudaStreamSynchronize between levels: 51 % of a 1 Mi-triangle commit and
// ~55 us per level even for a level of ONE node (profiles/r03a_build_1Mi_launches.csv).  Here EIGHT LANES share a
// node — lane k holds child k of the node being assembled, everything lives in registers, the greedy choices
// (open the largest child; give the best (child, octant) pair its slot) are 8-lane shuffle arg-max reductions — and
// the levels are separated by grid-wide barriers inside the kernel, so a commit never talks to the host.  The
// arithmetic is collapse_select's / encode_node8's (core.cuh, what tests/hostemu runs on the CPU), with the same
// tie rules (lowest child index, then lowest slot), so the tree is the same tree.
// ------------------------------------------------------------------------------------------------
struct CollapseArgs {
  int n;
  int* queue;                    // queue[j] = binary root of 8-wide node j
  unsigned int* counters;        // [0] nodes allocated, [1] triangles emitted, [2] levels, [3] nodes (final), [4] overflow
  const int4* info;              // refit_kernel's record per internal node
  const int* bfirst;
  const float *blo, *bhi, *llo, *lhi;
  const uint32_t* sorted;        // sorted position -> primID
  uint4* nodes;
  uint32_t node_cap;
  uint32_t* tri_prim;            // leaf-order slot -> primID (tri records are written by tri_records_kernel)
};

__device__ __forceinline__ float g8_max(unsigned m, float v) {
  v = fmaxf(v, __shfl_xor_sync(m, v, 1, 8)); v = fmaxf(v, __shfl_xor_sync(m, v, 2, 8)); return fmaxf(v, __shfl_xor_sync(m, v, 4, 8));
}
__device__ __forceinline__ float g8_min(unsigned m, float v) {
  v = fminf(v, __shfl_xor_sync(m, v, 1, 8)); v = fminf(v, __shfl_xor_sync(m, v, 2, 8)); return fminf(v, __shfl_xor_sync(m, v, 4, 8));
}
__device__ __forceinline__ uint32_t g8_pack(unsigned m, uint32_t byte, int l8) {   // lanes 0-3: slots 0..3, lanes 4-7: slots 4..7
  uint32_t w = byte << (8 * (l8 & 3));
  w |= __shfl_xor_sync(m, w, 1, 8);
  w |= __shfl_xor_sync(m, w, 2, 8);
  return w;
}

__device__ __forceinline__ void collapse_node8(const CollapseArgs& A, int j, unsigned gmask, int gshift, int l8) {
  const int b = A.queue[j];
  const int4 rb = A.info[b];
  int ref = 0, cnt = 0, cl = 0, cr = 0;
  float area = 0.f;
  auto load_child = [&](int c) {
    ref = c;
    if (c < 0) { cnt = 1; area = 0.f; cl = cr = 0; }
    else { const int4 ci = A.info[c]; cl = ci.x; cr = ci.y; cnt = ci.z; area = __int_as_float(ci.w); }
  };
  int nc = 2;
  if (l8 == 0) load_child(rb.x);
  else if (l8 == 1) load_child(rb.y);
  // phase 0: open the largest child that is not a leaf yet; phase 1 (SPRAY_SPLIT_LEAVES): spare slots split leaves
  for (int phase = 0; phase < (SPRAY_SPLIT_LEAVES ? 2 : 1); ++phase) {
    while (nc < 8) {
      const bool cand = l8 < nc && ref >= 0 && (phase == 0 ? cnt > SPRAY_LEAF_MAX : cnt <= SPRAY_LEAF_MAX) && area > -1.0f;
      const float key = cand ? area : -2.0f;
      const float m = g8_max(gmask, key);
      const unsigned bal = (__ballot_sync(gmask, cand && key == m) >> gshift) & 0xffu;
      if (!bal) break;
      const int best = __ffs((int)bal) - 1;
      const int L = __shfl_sync(gmask, cl, best, 8), R = __shfl_sync(gmask, cr, best, 8);
      if (l8 == best) load_child(L);
      else if (l8 == nc) load_child(R);
      ++nc;
    }
  }
  // child boxes, the box around them (for the octant costs only: the node's own box is the binary node's)
  const bool used = l8 < nc;
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  if (used) {
    const float* pl = ref < 0 ? A.llo + 3 * (size_t)(~ref) : A.blo + 3 * (size_t)ref;
    const float* ph = ref < 0 ? A.lhi + 3 * (size_t)(~ref) : A.bhi + 3 * (size_t)ref;
#pragma unroll
    for (int a = 0; a < 3; ++a) { lo[a] = pl[a]; hi[a] = ph[a]; }
  }
  float cost[8];
  {
    float cen[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float nlo = g8_min(gmask, lo[a]), nhi = g8_max(gmask, hi[a]);
      cen[a] = 0.5f * (lo[a] + hi[a]) - 0.5f * (nlo + nhi);
    }
#pragma unroll
    for (int s = 0; s < 8; ++s)
      cost[s] = ((s & 1) ? cen[0] : -cen[0]) + ((s & 2) ? cen[1] : -cen[1]) + ((s & 4) ? cen[2] : -cen[2]);
  }
  // greedy octant slots: the best remaining (child, slot) pair first; ties -> lowest child, then lowest slot
  int myslot = -1;
  unsigned freeslots = 0xffu;
  for (int it = 0; it < nc; ++it) {
    float bc = -INFINITY;
    int bs = -1;
    if (used && myslot < 0) {
#pragma unroll
      for (int s = 0; s < 8; ++s)
        if (((freeslots >> s) & 1u) && cost[s] > bc) { bc = cost[s]; bs = s; }
    }
    const float m = g8_max(gmask, bs >= 0 ? bc : -INFINITY);
    const unsigned bal = (__ballot_sync(gmask, bs >= 0 && bc == m) >> gshift) & 0xffu;
    int bk, bsel;
    if (bal) {
      bk = __ffs((int)bal) - 1;
      bsel = __shfl_sync(gmask, bs, bk, 8);
    } else {
      // NaN / -inf costs never win a comparison: first unplaced child -> first free slot (collapse_select's fallback)
      const unsigned un = (__ballot_sync(gmask, used && myslot < 0) >> gshift) & 0xffu;
      bk = __ffs((int)un) - 1;
      bsel = __ffs((int)freeslots) - 1;
    }
    if (l8 == bk) myslot = bsel;
    freeslots &= ~(1u << bsel);
  }
  // from here on lane s IS slot s
  int src = -1;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int ms = __shfl_sync(gmask, myslot, k, 8);
    if (ms == l8) src = k;
  }
  const bool has = src >= 0;
  const int sl = has ? src : 0;
  const int sref = __shfl_sync(gmask, ref, sl, 8), scnt = __shfl_sync(gmask, cnt, sl, 8);
  float slo[3], shi[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) { slo[a] = __shfl_sync(gmask, lo[a], sl, 8); shi[a] = __shfl_sync(gmask, hi[a], sl, 8); }
  const bool inner = has && sref >= 0 && scnt > SPRAY_LEAF_MAX;
  const bool leaf = has && !inner;
  const unsigned ib = (__ballot_sync(gmask, inner) >> gshift) & 0xffu;
  const int n_inner = __popc(ib), inner_idx = __popc(ib & ((1u << l8) - 1u));
  const int c = leaf ? scnt : 0;
  int incl = c;
#pragma unroll
  for (int off = 1; off < 8; off <<= 1) {
    const int t = __shfl_up_sync(gmask, incl, off, 8);
    if (l8 >= off) incl += t;
  }
  const int excl = incl - c, n_tris = __shfl_sync(gmask, incl, 7, 8);
  uint32_t meta = 0;
  if (inner) meta = (1u << 5) | (24u + (uint32_t)l8);
  else if (leaf) meta = ((c == 1 ? 1u : (c == 2 ? 3u : 7u)) << 5) | (uint32_t)excl;
  uint32_t child_base = 0, tri_base = 0;
  if (l8 == 0) {
    if (n_inner) child_base = atomicAdd(&A.counters[0], (unsigned)n_inner);
    if (n_tris) tri_base = atomicAdd(&A.counters[1], (unsigned)n_tris);
  }
  child_base = __shfl_sync(gmask, child_base, 0, 8);
  tri_base = __shfl_sync(gmask, tri_base, 0, 8);
  const bool fits = child_base + (uint32_t)n_inner <= A.node_cap && tri_base + (uint32_t)n_tris <= (uint32_t)A.n;
  if (!fits) { if (l8 == 0) A.counters[4] = 1u; }
  else if (inner) A.queue[child_base + inner_idx] = sref;
  else if (leaf) {
    const int first = sref < 0 ? ~sref : A.bfirst[sref];
    for (int k = 0; k < c; ++k) A.tri_prim[tri_base + excl + k] = A.sorted[first + k];
  }
  // quantise against the binary node's box: per axis the smallest exponent every child fits with its margin
  const float* nbl = A.blo + 3 * (size_t)b;
  const float* nbh = A.bhi + 3 * (size_t)b;
  uint32_t ql[3], qh[3], ebits[3];
  float porg[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float nlo = nbl[a], nhi = nbh[a];
    int e = quant_exp_init(nlo, nhi);
    for (;; ++e) {
      const float scale = u2f((uint32_t)(e + 127) << 23);
      const float p = nlo - scale;
      bool ok = true;
      uint32_t l = 255u, h = 0u;
      if (has) ok = quant_child(slo[a], shi[a], p, scale, &l, &h);
      const bool allok = ((__ballot_sync(gmask, !ok) >> gshift) & 0xffu) == 0u;
      if (allok || e >= 120) { ql[a] = l; qh[a] = h; ebits[a] = (uint32_t)(e + 127); porg[a] = p; break; }
    }
  }
  uint32_t w[7];
  w[0] = g8_pack(gmask, meta, l8);
#pragma unroll
  for (int a = 0; a < 3; ++a) { w[1 + a] = g8_pack(gmask, ql[a], l8); w[4 + a] = g8_pack(gmask, qh[a], l8); }
  uint32_t wh[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) wh[k] = __shfl_sync(gmask, w[k], 4, 8);
  if (l8 == 0 && (uint32_t)j < A.node_cap) {
    uint4* out = A.nodes + 5 * (size_t)j;
    out[0] = make_uint4(f2u(porg[0]), f2u(porg[1]), f2u(porg[2]), ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | (ib << 24));
    out[1] = make_uint4(child_base, tri_base, w[0], wh[0]);
    out[2] = make_uint4(w[1], wh[1], w[2], wh[2]);
    out[3] = make_uint4(w[3], wh[3], w[4], wh[4]);
    out[4] = make_uint4(w[5], wh[5], w[6], wh[6]);
  }
}

__global__ void __launch_bounds__(256) collapse_coop_kernel(CollapseArgs A) {
  cg::grid_group grid = cg::this_grid();
  const int l8 = threadIdx.x & 7;
  const int gshift = threadIdx.x & 24;
  const unsigned gmask = 0xffu << gshift;
  const size_t ngroups = (size_t)gridDim.x * (blockDim.x >> 3);
  const size_t gid = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  unsigned int begin = 0, end = 1, levels = 0;
  while (begin < end) {
    ++levels;
    for (size_t j = begin + gid; j < end; j += ngroups) collapse_node8(A, (int)j, gmask, gshift, l8);
    grid.sync();
    begin = end;
    end = min(*(volatile unsigned int*)&A.counters[0], A.node_cap);
    if (levels > 4096u) break;     // cannot happen (every level adds nodes or ends); keeps a corrupted tree from hanging the GPU
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) { A.counters[2] = levels; A.counters[3] = end; }
}

// 48-byte triangle records in BVH leaf order from the leaf-order primID list
__global__ void tri_records_kernel(const float* __restrict__ verts, const uint32_t* __restrict__ faces,
                                   const uint32_t* __restrict__ tri_prim, int n, float4* __restrict__ tris) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) write_tri(verts, faces, tri_prim[i], tris + 3 * (size_t)i);
}

// n == 1: a root with a single one-triangle leaf
__global__ void single_tri_kernel(const float* __restrict__ verts, const uint32_t* __restrict__ faces,
                                  const float* __restrict__ tlo, const float* __restrict__ thi,
                                  uint4* __restrict__ nodes, float4* __restrict__ tris) {
  CollapseOut co;
  co.n_inner = 0; co.n_tris = 1; co.n_leaf = 1; co.imask = 0;
  for (int s = 0; s < 8; ++s) co.meta[s] = 0;
  co.meta[0] = (uint8_t)(1u << 5);
  co.tri_first[0] = 0; co.tri_count[0] = 1;
  for (int a = 0; a < 3; ++a) { co.lo[0][a] = tlo[a]; co.hi[0][a] = thi[a]; }
  u4 q[5];
  encode_node8(tlo, thi, co, 0u, 0u, q);
  for (int k = 0; k < 5; ++k) nodes[k] = make_uint4(q[k].x, q[k].y, q[k].z, q[k].w);
  write_tri(verts, faces, 0u, tris);
}

// attribute record per primID: (n_k.xyz, rgb_k) for the triangle's three vertices (a3)
__global__ void attr_kernel(const uint32_t* __restrict__ faces, const float* __restrict__ normals,
                            const uint32_t* __restrict__ rgb, int nf, float4* __restrict__ attr) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nf) return;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const size_t v = faces[3 * (size_t)i + k];
    attr[3 * (size_t)i + k] =
        make_float4(normals[3 * v], normals[3 * v + 1], normals[3 * v + 2], __uint_as_float(rgb ? rgb[v] : 0u));
  }
}

__global__ void fill_u32_kernel(uint32_t* p, uint32_t v, size_t n) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// largest vertex index of the face list (spraygpu_scene_set_mesh validates it against nv: an out-of-range index of
// a malformed / truncated PLY would otherwise become out-of-bounds device WRITES in vcount_kernel at commit)
__global__ void __launch_bounds__(256) max_index_kernel(const uint32_t* __restrict__ idx, size_t n, uint32_t* __restrict__ out) {
  uint32_t m = 0u;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) m = max(m, idx[i]);
  m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

int launch_max_index(Device* dev, const uint32_t* d_idx, size_t n, uint32_t* d_out) {
  SG_CUDA(cudaMemsetAsync(d_out, 0, sizeof(uint32_t), upload_stream(dev)));
  const int grid = (int)std::min<size_t>((n + 255) / 256, (size_t)dev->sm_count * 8);
  max_index_kernel<<<grid, 256, 0, upload_stream(dev)>>>(d_idx, n, d_out);
  dev->launches++;
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

// ------------------------------------------------------------------------------------------------
int ensure_scratch(Scratch& s, size_t bytes) {
  if (bytes <= s.bytes) return SPRAYGPU_OK;
  if (s.ptr) cudaFree(s.ptr);
  s.ptr = nullptr; s.bytes = 0;
  size_t want = bytes + bytes / 4;
  SG_CUDA(cudaMalloc(&s.ptr, want));
  s.bytes = want;
  return SPRAYGPU_OK;
}

struct Carver {
  char* base; size_t off = 0;
  template <typename T> T* take(size_t n) {
    off = (off + 255) & ~(size_t)255;
    T* p = base ? (T*)(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
};

static inline int grid_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }

// SPRAY_BUILD_GRAPHS=0 disables the CUDA-graph replay of repeated commits (A/B knob)
static bool build_graphs_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SPRAY_BUILD_GRAPHS"); v = (e && *e == '0') ? 0 : 1; }
  return v == 1;
}

// one thread: the collapse's counters and root entry, the bounds accumulators (no host-staged copies in a commit)
__global__ void build_init_kernel(uint32_t* bounds, unsigned int* counters, int* queue, unsigned int single) {
  for (int a = 0; a < 3; ++a) {
    bounds[a] = bounds[6 + a] = 0xFFFFFFFFu;   // min slots
    bounds[3 + a] = bounds[9 + a] = 0u;        // max slots
  }
  counters[0] = 1u; counters[1] = single; counters[2] = 1u; counters[3] = 1u; counters[4] = 0u;   // single: n == 1, no collapse
  queue[0] = 0;
}

// Enqueues the whole build on the device's build stream and returns: no host round trip.  The few words the host
// wants back (node count, BVH8 depth, overflow flag, geometry bounds) land in the scene's pinned result record;
// scene_finalize() waits for them the first time somebody needs them.
int build_scene(Scene* sc) {
  Device* dev = sc->dev;
  cudaStream_t st = build_stream(dev);
  Scratch* scratch = &dev->scratch;
  const size_t nv = sc->nv, nf = sc->nf;
  SG_CHECK(nf >= 1 && nv >= 1, SPRAYGPU_ERR_STATE, "commit: empty mesh (nv=%zu nf=%zu)", nv, nf);
  if (nf <= SPRAY_POOL_MAX_TRIS && !dev->loading_async) {
    // small domain: one of the pool streams (round robin), ordered behind everything already on the device's stream
    // (a re-commit must not overwrite buffers a running trace still reads)
    if (sc->pool_slot < 0) sc->pool_slot = (int)(dev->pool_rr++ % SPRAY_BUILD_POOL);   // sticky: the graph of a re-commit names the stream
    const unsigned slot = (unsigned)sc->pool_slot;
    if (!dev->pool_stream[slot]) SG_CUDA(cudaStreamCreateWithFlags(&dev->pool_stream[slot], cudaStreamNonBlocking));
    if (!dev->pool_order) SG_CUDA(cudaEventCreateWithFlags(&dev->pool_order, cudaEventDisableTiming));
    SG_CUDA(cudaEventRecord(dev->pool_order, dev->stream));
    SG_CUDA(cudaStreamWaitEvent(dev->pool_stream[slot], dev->pool_order, 0));
    st = dev->pool_stream[slot];
    scratch = &dev->pool_scratch[slot];
  }
  SG_CHECK(nf <= SPRAYGPU_MAX_PRIMS, SPRAYGPU_ERR_LIMIT, "commit: %zu faces exceed the 23-bit primID field", nf);
  const int n = (int)nf;
  const int npairs = 3 * n;
  if (sc->pending) { int frc = scene_finalize(sc); if (frc) return frc; }

  if (!sc->ev0) {
    SG_CUDA(cudaEventCreate(&sc->ev0));
    SG_CUDA(cudaEventCreate(&sc->ev1));
    SG_CUDA(cudaMallocHost((void**)&sc->h_result, sizeof(uint32_t) * 32));
  }
  if (sc->ev_upload) SG_CUDA(cudaStreamWaitEvent(st, sc->ev_upload, 0));   // the mesh may have come in on another stream
  SG_CUDA(cudaEventRecord(sc->ev0, st));

  // ---- persistent outputs
  int grc;
  if (sc->has_xf) {
    if ((grc = grow_buffer(&sc->d_verts_xf, &sc->cap_verts_xf, sizeof(float) * 3 * nv))) return grc;
    sc->d_verts = sc->d_verts_xf;
  } else sc->d_verts = sc->d_verts_obj;
  if ((grc = grow_buffer(&sc->d_normals, &sc->cap_normals, sizeof(float) * 3 * nv))) return grc;
  if ((grc = grow_buffer(&sc->d_tris, &sc->cap_tris, sizeof(float4) * 3 * nf))) return grc;
  if ((grc = grow_buffer(&sc->d_attr, &sc->cap_attr, sizeof(float4) * 3 * nf))) return grc;
  // every 8-wide node but the root covers >= 4 triangles and a full node with k inner children has 8 - k leaf
  // children: N(c) <= (c - 1) / 3 nodes for c >= 4 triangles (induction over the collapse); measured 0.13-0.18 n
  const uint32_t node_cap = (uint32_t)(n / 3 + 2);
  if ((grc = grow_buffer(&sc->d_nodes, &sc->cap_nodes, sizeof(uint4) * 5 * (size_t)node_cap))) return grc;
  sc->node_cap = node_cap;

  // ---- scratch carve (two passes: size, then pointers)
  const int items_min = std::min(rs_items_for(npairs, dev->sm_count), rs_items_for(n, dev->sm_count));
  const int nblocks_max = (npairs + RS_THREADS * items_min - 1) / (RS_THREADS * items_min);
  uint32_t *kA, *kB, *vA, *vB, *hist, *vstart, *vend, *vcnt_buf, *bounds, *tri_prim;
  unsigned int* counters;
  float *tlo, *thi, *llo, *lhi, *blo, *bhi;
  int *bl, *br, *bfirst, *blast, *parent, *flags, *queue;
  int4* info;
  auto carve = [&](Carver& c) {
    kA = c.take<uint32_t>(npairs); kB = c.take<uint32_t>(npairs);
    vA = c.take<uint32_t>(npairs); vB = c.take<uint32_t>(npairs);
    hist = c.take<uint32_t>((size_t)256 * nblocks_max + 1024);
    vstart = c.take<uint32_t>(nv); vend = c.take<uint32_t>(nv); vcnt_buf = c.take<uint32_t>(nv);
    bounds = c.take<uint32_t>(12); counters = c.take<unsigned int>(8);
    tlo = c.take<float>(3 * nf); thi = c.take<float>(3 * nf);
    llo = c.take<float>(3 * nf); lhi = c.take<float>(3 * nf);
    blo = c.take<float>(3 * nf); bhi = c.take<float>(3 * nf);
    bl = c.take<int>(nf); br = c.take<int>(nf); bfirst = c.take<int>(nf); blast = c.take<int>(nf);
    parent = c.take<int>(2 * nf); flags = c.take<int>(nf); queue = c.take<int>(node_cap);
    info = c.take<int4>(nf); tri_prim = c.take<uint32_t>(nf);
  };
  Carver sizer{nullptr};
  carve(sizer);
  int rc = ensure_scratch(*scratch, sizer.off + 256);
  if (rc) return rc;
  Carver carver{(char*)scratch->ptr};
  carve(carver);

  // Everything from here to the result copies is ~30 launches of fixed shape for a given (mesh size, buffers, stream):
  // the second commit with the same shape captures them into a CUDA graph, later ones replay it with ONE launch.  A
  // re-committed cache block (out-of-core LRU, the reference's rtcCommit on every miss) and the many small domains of
  // a C2 / C5 scene are enqueue-bound otherwise: ~35 API calls = 0.14 ms of host time per commit (r04x).
  auto enqueue = [&]() -> int {
    const int T = 256;
    build_init_kernel<<<1, 1, 0, st>>>(bounds, counters, queue, n == 1 ? 1u : 0u);
    dev->launches++;
    // ---- transform
    if (sc->has_xf) {
      Mat4 M; memcpy(M.m, sc->xf, sizeof(M.m));
      transform_kernel<<<grid_for(nv, T), T, 0, st>>>(sc->d_verts_obj, sc->d_verts, nv, M);
      dev->launches++;
    }
    // ---- normals: valence count -> segments -> face lists -> per-vertex sorted sums (no global sort)
    SG_CUDA(cudaMemsetAsync(vend, 0, sizeof(uint32_t) * nv, st));        // vend = per-vertex fill cursor
    SG_CUDA(cudaMemsetAsync(vcnt_buf, 0, sizeof(uint32_t) * nv, st));    // valence
    SG_CUDA(cudaMemsetAsync(counters + 5, 0, sizeof(unsigned int) * 2, st));   // [5] segment cursor, [6] heavy vertices
    {
      uint32_t* vcnt = vcnt_buf;
      uint32_t* vlist = kA;                 // npairs face ids, grouped by vertex
      uint32_t* heavy = vB;                 // vertices with more than NRM_LIGHT_MAX faces (at most npairs / 33)
      vcount_kernel<<<grid_for(npairs, T), T, 0, st>>>(sc->d_faces, npairs, vcnt);
      valloc_kernel<<<grid_for(nv, 1024), 1024, 0, st>>>(vcnt, nv, vstart, counters + 5);
      vfill_kernel<<<grid_for(npairs, T), T, 0, st>>>(sc->d_faces, npairs, vstart, vend, vlist);
      normals_light_kernel<<<grid_for(nv, T), T, 0, st>>>(sc->d_verts, sc->d_faces, vlist, vstart, vcnt, sc->d_normals, nv, heavy,
                                                          counters + 6);
      normals_heavy_kernel<<<std::min(dev->sm_count * 4, npairs / (NRM_LIGHT_MAX + 1) + 1), 256, 0, st>>>(
          sc->d_verts, sc->d_faces, vlist, vstart, vcnt, sc->d_normals, heavy, counters + 6);
      attr_kernel<<<grid_for(nf, T), T, 0, st>>>(sc->d_faces, sc->d_normals, sc->d_rgb, n, sc->d_attr);
      dev->launches += 6;
    }

    // ---- triangle boxes, Morton keys, sort
    tri_bounds_kernel<<<std::min(grid_for(nf, 256), dev->sm_count * 8), 256, 0, st>>>(sc->d_verts, sc->d_faces, n, tlo, thi, bounds);
    dev->launches++;
    if (n == 1) {
      single_tri_kernel<<<1, 1, 0, st>>>(sc->d_verts, sc->d_faces, tlo, thi, sc->d_nodes, sc->d_tris);
      dev->launches++;
    } else {
      morton_kernel<<<grid_for(nf, T), T, 0, st>>>(tlo, thi, n, bounds, kA, vA);
      dev->launches++;
      rc = radix_sort(dev, st, &kA, &vA, &kB, &vB, hist, n, 30);
      if (rc) return rc;
      leaf_boxes_kernel<<<grid_for(nf, T), T, 0, st>>>(vA, tlo, thi, n, llo, lhi);
      lbvh_kernel<<<grid_for(nf, T), T, 0, st>>>(kA, n, bl, br, bfirst, blast, parent);
      SG_CUDA(cudaMemsetAsync(flags, 0, sizeof(int) * nf, st));
      refit_kernel<<<grid_for(nf, T), T, 0, st>>>(parent, bl, br, bfirst, blast, llo, lhi, blo, bhi, info, flags, n);
      dev->launches += 3;
      // ---- collapse: one cooperative launch walks the levels (queue[j] = binary root of 8-wide node j)
      if (dev->collapse_blocks_per_sm == 0) {
        int bps = 0;
        SG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, collapse_coop_kernel, 256, 0));
        SG_CHECK(bps > 0, SPRAYGPU_ERR_LIMIT, "commit: the collapse kernel does not fit an SM");
        dev->collapse_blocks_per_sm = bps;
      }
      CollapseArgs A;
      A.n = n; A.queue = queue; A.counters = counters; A.info = info; A.bfirst = bfirst;
      A.blo = blo; A.bhi = bhi; A.llo = llo; A.lhi = lhi; A.sorted = vA; A.nodes = sc->d_nodes; A.node_cap = node_cap;
      A.tri_prim = tri_prim;
      // a level is at most node_cap nodes wide, 32 nodes per 256-thread CTA
      const int grid = std::max(1, std::min(dev->collapse_blocks_per_sm * dev->sm_count, (int)((node_cap + 31) / 32)));
      void* args[] = {(void*)&A};
      SG_CUDA(cudaLaunchCooperativeKernel((const void*)collapse_coop_kernel, dim3(grid), dim3(256), args, 0, st));
      tri_records_kernel<<<grid_for(nf, T), T, 0, st>>>(sc->d_verts, sc->d_faces, tri_prim, n, sc->d_tris);
      dev->launches += 2;
    }
    // ---- results: 8 counter words + 12 bounds words into the scene's pinned record
    SG_CUDA(cudaMemcpyAsync(sc->h_result, counters, sizeof(uint32_t) * 8, cudaMemcpyDeviceToHost, st));
    SG_CUDA(cudaMemcpyAsync(sc->h_result + 8, bounds, sizeof(uint32_t) * 12, cudaMemcpyDeviceToHost, st));
    return SPRAYGPU_OK;
  };
  uint64_t key = 1469598103934665603ull;
  {
    const void* ptrs[] = {sc->d_verts_obj, sc->d_faces, sc->d_rgb, sc->d_verts, sc->d_normals, sc->d_tris, sc->d_attr, sc->d_nodes,
                          scratch->ptr, (const void*)st, sc->h_result};
    auto mix = [&](const void* p, size_t nbytes) { const unsigned char* c = (const unsigned char*)p; for (size_t i = 0; i < nbytes; ++i) { key ^= c[i]; key *= 1099511628211ull; } };
    mix(ptrs, sizeof(ptrs)); mix(&nv, sizeof(nv)); mix(&nf, sizeof(nf)); mix(&sc->has_xf, sizeof(sc->has_xf));
    if (sc->has_xf) mix(sc->xf, sizeof(sc->xf));
  }
  bool launched = false;
  if (build_graphs_enabled() && !sc->graph_failed) {
    if (sc->graph_exec && sc->graph_key == key) {
      SG_CUDA(cudaGraphLaunch(sc->graph_exec, st));
      dev->launches += sc->graph_launches;
      launched = true;
    } else if (sc->seen_key == key) {
      if (sc->graph_exec) { cudaGraphExecDestroy(sc->graph_exec); sc->graph_exec = nullptr; }
      if (cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) == cudaSuccess) {
        const uint64_t l0 = dev->launches;
        const int erc = enqueue();
        cudaGraph_t g = nullptr;
        const cudaError_t ce = cudaStreamEndCapture(st, &g);
        if (erc == SPRAYGPU_OK && ce == cudaSuccess && g && cudaGraphInstantiate(&sc->graph_exec, g, 0) == cudaSuccess) {
          sc->graph_key = key;
          sc->graph_launches = (uint32_t)(dev->launches - l0);
          SG_CUDA(cudaGraphLaunch(sc->graph_exec, st));
          launched = true;
        } else {
          sc->graph_failed = true;        // whatever this driver cannot capture: plain launches from now on
          sc->graph_exec = nullptr;
          dev->launches = l0;
          cudaGetLastError();
        }
        if (g) cudaGraphDestroy(g);
      } else cudaGetLastError();
    }
  }
  sc->seen_key = key;
  if (!launched) { const int erc = enqueue(); if (erc) return erc; }
  SG_CUDA(cudaEventRecord(sc->ev1, st));
  SG_CUDA(cudaGetLastError());
  sc->pending = true;
  sc->committed = true;
  return SPRAYGPU_OK;
}

// waits for an enqueued build and takes its results; the checks a commit owes its caller happen here
int scene_finalize(Scene* sc) {
  if (!sc->pending) return SPRAYGPU_OK;
  SG_CUDA(cudaEventSynchronize(sc->ev1));
  sc->pending = false;
  const uint32_t* h = sc->h_result;
  sc->n_nodes = h[3];
  sc->bvh_depth = h[2];
  for (int a = 0; a < 6; ++a) {
    const uint32_t s = h[8 + 6 + a];
    const uint32_t u = (s & 0x80000000u) ? (s & 0x7fffffffu) : ~s;
    memcpy(&sc->bounds[a], &u, 4);
  }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, sc->ev0, sc->ev1);
  sc->build_ms = ms;
  if (h[4] != 0u || h[1] != (uint32_t)sc->nf || sc->n_nodes > sc->node_cap) {
    sc->committed = false;
    set_error("commit: the collapse overflowed its buffers (%u nodes of %u, %u of %zu triangles)", h[0], sc->node_cap, h[1], sc->nf);
    return SPRAYGPU_ERR_STATE;
  }
  // the traversal stack holds one entry per BVH8 level (core.cuh, SPRAY_STACK_SIZE): refuse deeper trees here
  // instead of dropping subtrees at trace time
  if (sc->bvh_depth > (uint32_t)SPRAY_STACK_SIZE) {
    sc->committed = false;
    set_error("commit: BVH8 depth %u exceeds the traversal stack (%d)", sc->bvh_depth, SPRAY_STACK_SIZE);
    return SPRAYGPU_ERR_LIMIT;
  }
  return SPRAYGPU_OK;
}


// ------------------------------------------------------------------------------------------------
// spraygpu_compute_normals_static — the per-vertex normals stage for hosts that must not allocate.
//
// SpRay's in-situ mode runs the renderer INSIDE a simulation.  Such a host has usually handed all of HBM to its own
// pool at start-up, so a cudaMalloc from a plug-in either fails or takes memory the solver counted on; and cudaMalloc /
// cudaFree synchronise the whole device, which stalls the solver's streams in the middle of a time step.  The regular
// path (spraygpu_scene_commit) grows a scratch buffer with cudaMalloc.  This entry point runs the same five kernels
// (vcount -> valloc -> vfill -> normals_light -> normals_heavy, unchanged) with their work space carved out of the
// file-scope device array below instead: valence counts, segment starts, fill cursors, the per-vertex face lists and
// the list of heavy vertices.  No cudaMalloc, cudaFree, event / stream creation or other allocating call is made.
//
// The array is part of the library image: every GPU that loads libspraygpu holds its ~4.6 MB until the process exits,
// and all calls on one GPU share it.  Nothing is kept between calls on purpose — the counters, the valences and the fill
// cursors are zeroed at the start of every call because the previous call's values are still there — but two calls whose
// kernels overlap on the GPU (different streams, or different host threads) trample each other's lists.
// ------------------------------------------------------------------------------------------------
constexpr size_t NS_MAXV = SPRAYGPU_STATIC_MAX_VERTS, NS_MAXF = SPRAYGPU_STATIC_MAX_FACES;
constexpr size_t ns_pad(size_t words) { return (words + 63) & ~(size_t)63; }      // every region starts 256-byte aligned
constexpr size_t NS_VCNT = 0;                                   // [nv]  valence
constexpr size_t NS_VSTART = NS_VCNT + ns_pad(NS_MAXV);         // [nv]  first slot of the vertex's face list
constexpr size_t NS_VFILL = NS_VSTART + ns_pad(NS_MAXV);        // [nv]  fill cursor                         (starts at 0)
constexpr size_t NS_VLIST = NS_VFILL + ns_pad(NS_MAXV);         // [3 nf] face ids grouped by vertex
constexpr size_t NS_HEAVY = NS_VLIST + ns_pad(3 * NS_MAXF);     // vertices with > NRM_LIGHT_MAX faces: < 3 nf / 32
constexpr size_t NS_CTR = NS_HEAVY + ns_pad(3 * NS_MAXF / NRM_LIGHT_MAX + 1);   // [0] segment cursor [1] heavy count [2] max index
constexpr size_t NS_WORDS = NS_CTR + 64;
static __device__ __align__(256) uint32_t g_normals_scratch[NS_WORDS];

// In HOST code `g_normals_scratch` is only the host-side shadow of the symbol: handing it to a kernel would fault (or,
// on a system where the GPU can reach host memory, quietly work on the wrong bytes).  The device address comes from
// cudaGetSymbolAddress, per device (every GPU has its own copy), asked for with that device current.
constexpr int NS_MAX_DEVICES = 64;
static uint32_t* g_ns_addr[NS_MAX_DEVICES];
static int g_ns_sms[NS_MAX_DEVICES];

static int normals_static_setup(int device, uint32_t** base, int* sm_count) {
  SG_CHECK(device >= 0 && device < NS_MAX_DEVICES, SPRAYGPU_ERR_LIMIT, "compute_normals_static: CUDA device %d (at most %d)",
           device, NS_MAX_DEVICES);
  uint32_t* p = __atomic_load_n(&g_ns_addr[device], __ATOMIC_ACQUIRE);
  if (!p) {
    // racing first calls resolve the same address twice: harmless
    void* sym = nullptr;
    SG_CUDA(cudaGetSymbolAddress(&sym, g_normals_scratch));
    SG_CHECK(sym != nullptr, SPRAYGPU_ERR_CUDA, "compute_normals_static: no device address for the scratch array");
    int sms = 0;
    SG_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    g_ns_sms[device] = sms > 0 ? sms : 148;
    p = (uint32_t*)sym;
    __atomic_store_n(&g_ns_addr[device], p, __ATOMIC_RELEASE);
  }
  *base = p;
  *sm_count = g_ns_sms[device];
  return SPRAYGPU_OK;
}

static int ns_device_pointer(const void* p, int device, const char* what) {
  cudaPointerAttributes at;
  SG_CUDA(cudaPointerGetAttributes(&at, p));
  SG_CHECK((at.type == cudaMemoryTypeDevice && at.device == device) || at.type == cudaMemoryTypeManaged, SPRAYGPU_ERR_ARG,
           "compute_normals_static: %s must be device memory of the current GPU (%d); host arrays would need a staging "
           "allocation, which this entry point never makes", what, device);
  return SPRAYGPU_OK;
}


// ------------------------------------------------------------------------------------------------
// Counter-based triangle soup (the synthetic C3 / C4 domains of the benchmark): every triangle is a pure
// function of (key, primID), so the bench scene is generated on the device in milliseconds AND can be handed to the
// CPU checker, which restates this function in C, bit for bit.  Philox-4x32-10 (Salmon et al., SC'11), key = (key, 0),
// counter = (primID, block, 0, 0).  Only IEEE-exact float operations are used (add, mul, div, sqrt, min, max; this file
// is built with -fmad=false), and the direction of an edge comes from rejection sampling in the cube — no sin / cos /
// log, whose device and libm versions differ in the last place.
//   block 0      : centre x, y, z (uniform in the cell), length factor of edge 1
//   block 1      : length factor of edge 2
//   block 2 + k  : k-th candidate direction of edge 1, k < 16 (accepted when 1e-4 < |d|^2 <= 1), then (1, 0, 0)
//   block 18 + k : the same for edge 2
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t k0) {
  uint32_t x0 = c0, x1 = c1, x2 = 0u, x3 = 0u, ka = k0, kb = 0u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t h0 = __umulhi(0xD2511F53u, x0), l0 = 0xD2511F53u * x0;
    const uint32_t h1 = __umulhi(0xCD9E8D57u, x2), l1 = 0xCD9E8D57u * x2;
    const uint32_t y0 = h1 ^ x1 ^ ka, y2 = h0 ^ x3 ^ kb;
    x0 = y0; x1 = l1; x2 = y2; x3 = l0;
    ka += 0x9E3779B9u; kb += 0xBB67AE85u;
  }
  return make_uint4(x0, x1, x2, x3);
}
__device__ __forceinline__ float soup_u01(uint32_t x) { return mul_rn((float)(x >> 8), 5.9604644775390625e-08f); }   // [0, 1), exact
__device__ __forceinline__ void soup_edge(uint32_t key, uint32_t prim, uint32_t block0, float len, float* ex, float* ey, float* ez) {
  float dx = 1.0f, dy = 0.0f, dz = 0.0f;
  for (uint32_t k = 0; k < 16u; ++k) {
    const uint4 r = philox4x32_10(prim, block0 + k, key);
    const float x = sub_rn(mul_rn(2.0f, soup_u01(r.x)), 1.0f), y = sub_rn(mul_rn(2.0f, soup_u01(r.y)), 1.0f),
                z = sub_rn(mul_rn(2.0f, soup_u01(r.z)), 1.0f);
    const float q = add_rn(add_rn(mul_rn(x, x), mul_rn(y, y)), mul_rn(z, z));
    if (q > 1.0e-4f && q <= 1.0f) {
      const float inv = div_rn(1.0f, sqrt_rn(q));
      dx = mul_rn(x, inv); dy = mul_rn(y, inv); dz = mul_rn(z, inv);
      break;
    }
  }
  *ex = mul_rn(dx, len); *ey = mul_rn(dy, len); *ez = mul_rn(dz, len);
}
struct SoupArgs { uint32_t key; uint32_t n; float lo[3], hi[3], h; };
__global__ void __launch_bounds__(256) soup_kernel(SoupArgs A, float* __restrict__ xyz, uint32_t* __restrict__ idx, uint32_t* __restrict__ rgb) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= A.n) return;
  const uint4 r0 = philox4x32_10(p, 0u, A.key), r1 = philox4x32_10(p, 1u, A.key);
  float c[3], e1[3], e2[3];
  c[0] = add_rn(A.lo[0], mul_rn(soup_u01(r0.x), sub_rn(A.hi[0], A.lo[0])));
  c[1] = add_rn(A.lo[1], mul_rn(soup_u01(r0.y), sub_rn(A.hi[1], A.lo[1])));
  c[2] = add_rn(A.lo[2], mul_rn(soup_u01(r0.z), sub_rn(A.hi[2], A.lo[2])));
  const float len1 = mul_rn(add_rn(0.5f, mul_rn(1.5f, soup_u01(r0.w))), A.h);
  const float len2 = mul_rn(add_rn(0.5f, mul_rn(1.5f, soup_u01(r1.x))), A.h);
  soup_edge(A.key, p, 2u, len1, &e1[0], &e1[1], &e1[2]);
  soup_edge(A.key, p, 18u, len2, &e2[0], &e2[1], &e2[2]);
  float* v = xyz + 9 * (size_t)p;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    v[a] = fminf(fmaxf(c[a], A.lo[a]), A.hi[a]);
    v[3 + a] = fminf(fmaxf(add_rn(c[a], e1[a]), A.lo[a]), A.hi[a]);
    v[6 + a] = fminf(fmaxf(add_rn(c[a], e2[a]), A.lo[a]), A.hi[a]);
  }
#pragma unroll
  for (uint32_t k = 0; k < 3u; ++k) {
    const uint32_t i = 3u * p + k;
    if (idx) idx[i] = i;
    if (rgb) {
      const uint32_t hsh = (i * 2654435761u) ^ (i >> 13);
      rgb[i] = ((64u + (hsh & 0x7Fu)) << 16) | ((64u + ((hsh >> 8) & 0x7Fu)) << 8) | (64u + ((hsh >> 16) & 0x7Fu));
    }
  }
}

}  // namespace spraygpu

using namespace spraygpu;

extern "C" int spraygpu_compute_normals_static(const float* xyz, size_t nv, const uint32_t* idx, size_t nf, float* normals,
                                               void* cuda_stream) {
  SG_CHECK(xyz && idx && normals && nv >= 1 && nf >= 1, SPRAYGPU_ERR_ARG, "compute_normals_static: bad argument (nv=%zu nf=%zu)", nv, nf);
  SG_CHECK(nv <= NS_MAXV && nf <= NS_MAXF, SPRAYGPU_ERR_LIMIT,
           "compute_normals_static: %zu vertices / %zu faces exceed the built-in scratch (at most %zu / %zu); use "
           "spraygpu_scene_commit, which allocates",
           nv, nf, NS_MAXV, NS_MAXF);
  int device = -1;
  SG_CUDA(cudaGetDevice(&device));
  uint32_t* S = nullptr;
  int sms = 148, rc;
  if ((rc = normals_static_setup(device, &S, &sms))) return rc;   // first call per GPU: resolves the array's device address
  if ((rc = ns_device_pointer(xyz, device, "xyz")) || (rc = ns_device_pointer(idx, device, "idx")) ||
      (rc = ns_device_pointer(normals, device, "normals")))
    return rc;
  cudaStream_t st = (cudaStream_t)cuda_stream;
  uint32_t *vcnt = S + NS_VCNT, *vstart = S + NS_VSTART, *vfill = S + NS_VFILL, *vlist = S + NS_VLIST, *heavy = S + NS_HEAVY;
  unsigned int* ctr = S + NS_CTR;
  const int npairs = (int)(3 * nf), T = 256;

  // what has to start at zero (the array still holds the previous call's values)
  SG_CUDA(cudaMemsetAsync(ctr, 0, sizeof(uint32_t) * 4, st));
  SG_CUDA(cudaMemsetAsync(vcnt, 0, sizeof(uint32_t) * nv, st));
  SG_CUDA(cudaMemsetAsync(vfill, 0, sizeof(uint32_t) * nv, st));

  // an index >= nv would become out-of-bounds writes into the scratch array: check before the scatter, as
  // spraygpu_scene_set_mesh does (4 bytes come back to the host; the call waits for the stream here)
  max_index_kernel<<<(int)std::min<size_t>(((size_t)npairs + 255) / 256, (size_t)sms * 8), 256, 0, st>>>(idx, (size_t)npairs, ctr + 2);
  SG_CUDA(cudaGetLastError());
  uint32_t h_max = 0;
  SG_CUDA(cudaMemcpyAsync(&h_max, ctr + 2, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  SG_CUDA(cudaStreamSynchronize(st));
  SG_CHECK((size_t)h_max < nv, SPRAYGPU_ERR_ARG, "compute_normals_static: face index %u out of range (nv = %zu)", h_max, nv);

  vcount_kernel<<<grid_for((size_t)npairs, T), T, 0, st>>>(idx, npairs, vcnt);
  valloc_kernel<<<grid_for(nv, 1024), 1024, 0, st>>>(vcnt, nv, vstart, ctr + 0);
  vfill_kernel<<<grid_for((size_t)npairs, T), T, 0, st>>>(idx, npairs, vstart, vfill, vlist);
  normals_light_kernel<<<grid_for(nv, T), T, 0, st>>>(xyz, idx, vlist, vstart, vcnt, normals, nv, heavy, ctr + 1);
  normals_heavy_kernel<<<std::min(sms * 4, npairs / (NRM_LIGHT_MAX + 1) + 1), 256, 0, st>>>(xyz, idx, vlist, vstart, vcnt, normals,
                                                                                          heavy, ctr + 1);
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}

extern "C" int spraygpu_gen_soup(uint32_t key, size_t ntris, const float lo[3], const float hi[3], float h, float* xyz, uint32_t* idx,
                                 uint32_t* rgb, void* cuda_stream) {
  SG_CHECK(lo && hi && xyz && ntris >= 1 && ntris <= SPRAYGPU_MAX_PRIMS, SPRAYGPU_ERR_ARG, "gen_soup: bad argument (ntris=%zu)", ntris);
  SoupArgs A;
  A.key = key; A.n = (uint32_t)ntris; A.h = h;
  for (int a = 0; a < 3; ++a) { A.lo[a] = lo[a]; A.hi[a] = hi[a]; }
  soup_kernel<<<grid_for(ntris, 256), 256, 0, (cudaStream_t)cuda_stream>>>(A, xyz, idx, rgb);
  SG_CUDA(cudaGetLastError());
  return SPRAYGPU_OK;
}
