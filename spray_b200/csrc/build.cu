// build.cu — per-domain acceleration-structure build on the GPU (= what rtcCommit does for the
// reference, src/render/trimesh_buffer.cc:189-225), plus the vertex transform (:141-157) and the
// per-vertex normals (:267-326).
//
// Pipeline (all on dev->stream, no host round trips except one 8-byte counter read per BVH8 level):
//   transform -> (vertex,face) pair sort -> deterministic per-vertex normal sums
//   -> triangle boxes + centroid bounds -> 30-bit Morton keys -> LSD radix sort
//   -> Karras LBVH -> bottom-up refit -> level-synchronous collapse into 80-byte 8-wide nodes
//   -> triangle records (48 B, leaf order) + attribute records (48 B, primID order)
#include <algorithm>

#include "internal.cuh"

namespace spraygpu {

// ------------------------------------------------------------------------------------------------
// LSD radix sort, 8 bits per pass, stable, key+value (uint32).  Tile = 256 threads x 16 keys.
// ------------------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;

__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint32_t* __restrict__ keys, int n, int shift,
                                                             uint32_t* __restrict__ hist, int nblocks) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int base = blockIdx.x * RS_TILE;
#pragma unroll
  for (int r = 0; r < RS_ITEMS; ++r) {
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
                      const uint32_t* __restrict__ sums) {
  __shared__ uint32_t digit_base[256];
  __shared__ uint32_t warp_cnt[RS_THREADS / 32][256];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  {
    const int idx = tid * nblocks + blockIdx.x;
    digit_base[tid] = hist[idx] + sums[idx / SC_CHUNK];
  }
  const int base = blockIdx.x * RS_TILE;
  for (int r = 0; r < RS_ITEMS; ++r) {
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
static int radix_sort(Device* dev, uint32_t** k, uint32_t** v, uint32_t** k2, uint32_t** v2, uint32_t* hist, int n,
                      int bits) {
  if (n <= 1) return SPRAYGPU_OK;
  const int nblocks = (n + RS_TILE - 1) / RS_TILE;
  const int passes = (bits + 7) / 8;
  const int total = 256 * nblocks, nchunks = (total + SC_CHUNK - 1) / SC_CHUNK;
  SG_CHECK(nchunks <= 1024, SPRAYGPU_ERR_LIMIT, "radix sort: %d keys exceed the two-level scan (max %d)", n, 1024 * SC_CHUNK / 256 * RS_TILE);
  uint32_t* sums = hist + (size_t)256 * nblocks;   // 1024 chunk totals live behind the counters
  for (int p = 0; p < passes; ++p) {
    const int shift = 8 * p;
    rs_hist_kernel<<<nblocks, RS_THREADS, 0, dev->stream>>>(*k, n, shift, hist, nblocks);
    rs_scan_chunks_kernel<<<nchunks, 1024, 0, dev->stream>>>(hist, total, sums);
    rs_scan_sums_kernel<<<1, 1024, 0, dev->stream>>>(sums, nchunks);
    rs_scatter_kernel<<<nblocks, RS_THREADS, 0, dev->stream>>>(*k, *v, *k2, *v2, n, shift, hist, nblocks, sums);
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

__global__ void pairs_kernel(const uint32_t* __restrict__ faces, uint32_t* __restrict__ keys,
                             uint32_t* __restrict__ vals, int npairs) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npairs) return;
  keys[i] = faces[i];
  vals[i] = (uint32_t)i / 3u;
}

__global__ void segments_kernel(const uint32_t* __restrict__ skeys, int npairs, uint32_t* __restrict__ vstart,
                                uint32_t* __restrict__ vend) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npairs) return;
  uint32_t k = skeys[i];
  if (i == 0 || skeys[i - 1] != k) vstart[k] = (uint32_t)i;
  if (i == npairs - 1 || skeys[i + 1] != k) vend[k] = (uint32_t)i + 1u;
}

// TriMeshBuffer::computeNormals, trimesh_buffer.cc:267-326.  The reference adds face normals to a
// vertex in increasing face order; the stable sort preserved that order, so each vertex's float sum is
// bit-identical to the serial loop.
__global__ void normals_kernel(const float* __restrict__ verts, const uint32_t* __restrict__ faces,
                               const uint32_t* __restrict__ sfaces, const uint32_t* __restrict__ vstart,
                               const uint32_t* __restrict__ vend, float* __restrict__ normals, size_t nv) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= nv) return;
  float nx = 0.0f, ny = 0.0f, nz = 0.0f;
  const uint32_t b = vstart[v], e = vend[v];
  for (uint32_t i = b; i < e; ++i) {
    const uint32_t f = sfaces[i];
    const uint32_t i0 = faces[3 * (size_t)f], i1 = faces[3 * (size_t)f + 1], i2 = faces[3 * (size_t)f + 2];
    vec3 v0 = mk3(verts[3 * (size_t)i0], verts[3 * (size_t)i0 + 1], verts[3 * (size_t)i0 + 2]);
    vec3 v1 = mk3(verts[3 * (size_t)i1], verts[3 * (size_t)i1 + 1], verts[3 * (size_t)i1 + 2]);
    vec3 v2 = mk3(verts[3 * (size_t)i2], verts[3 * (size_t)i2 + 1], verts[3 * (size_t)i2 + 2]);
    vec3 u = gsub(v1, v0), w = gsub(v2, v0);
    float fx = sub_rn(mul_rn(u.y, w.z), mul_rn(u.z, w.y));
    float fy = sub_rn(mul_rn(u.z, w.x), mul_rn(u.x, w.z));
    float fz = sub_rn(mul_rn(u.x, w.y), mul_rn(u.y, w.x));
    nx = add_rn(nx, fx); ny = add_rn(ny, fy); nz = add_rn(nz, fz);
  }
  normals[3 * v] = nx; normals[3 * v + 1] = ny; normals[3 * v + 2] = nz;
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

__global__ void refit_kernel(const int* __restrict__ parent, const int* __restrict__ bl, const int* __restrict__ br,
                             const float* __restrict__ llo, const float* __restrict__ lhi, float* blo, float* bhi,
                             int* flags, int n) {
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

// one thread = one 8-wide node of the current level
__global__ void __launch_bounds__(128)
    collapse_kernel(int begin, int end, int* __restrict__ queue, unsigned int* __restrict__ counters,
                    const int* __restrict__ bl, const int* __restrict__ br, const int* __restrict__ bfirst,
                    const int* __restrict__ blast, const float* __restrict__ blo, const float* __restrict__ bhi,
                    const float* __restrict__ llo, const float* __restrict__ lhi,
                    const uint32_t* __restrict__ sorted, const float* __restrict__ verts,
                    const uint32_t* __restrict__ faces, uint4* __restrict__ nodes, float4* __restrict__ tris) {
  int j = begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= end) return;
  const int b = queue[j];
  CollapseOut co;
  collapse_select(b, bl, br, bfirst, blast, blo, bhi, llo, lhi, &co);
  const uint32_t child_base = co.n_inner ? atomicAdd(&counters[0], (unsigned)co.n_inner) : 0u;
  const uint32_t tri_base = co.n_tris ? atomicAdd(&counters[1], (unsigned)co.n_tris) : 0u;
  for (int r = 0; r < co.n_inner; ++r) queue[child_base + r] = co.inner_ref[r];
  u4 q[5];
  encode_node8(blo + 3 * (size_t)b, bhi + 3 * (size_t)b, co, child_base, tri_base, q);
#pragma unroll
  for (int k = 0; k < 5; ++k) nodes[5 * (size_t)j + k] = make_uint4(q[k].x, q[k].y, q[k].z, q[k].w);
  uint32_t off = 0;
  for (int l = 0; l < co.n_leaf; ++l)
    for (int k = 0; k < co.tri_count[l]; ++k, ++off)
      write_tri(verts, faces, sorted[co.tri_first[l] + k], tris + 3 * (size_t)(tri_base + off));
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
// a malformed / truncated PLY would otherwise become out-of-bounds device WRITES in segments_kernel at commit)
__global__ void __launch_bounds__(256) max_index_kernel(const uint32_t* __restrict__ idx, size_t n, uint32_t* __restrict__ out) {
  uint32_t m = 0u;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) m = max(m, idx[i]);
  m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

int launch_max_index(Device* dev, const uint32_t* d_idx, size_t n, uint32_t* d_out) {
  SG_CUDA(cudaMemsetAsync(d_out, 0, sizeof(uint32_t), dev->stream));
  const int grid = (int)std::min<size_t>((n + 255) / 256, (size_t)dev->sm_count * 8);
  max_index_kernel<<<grid, 256, 0, dev->stream>>>(d_idx, n, d_out);
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

int build_scene(Scene* sc) {
  Device* dev = sc->dev;
  cudaStream_t st = dev->stream;
  const size_t nv = sc->nv, nf = sc->nf;
  SG_CHECK(nf >= 1 && nv >= 1, SPRAYGPU_ERR_STATE, "commit: empty mesh (nv=%zu nf=%zu)", nv, nf);
  SG_CHECK(nf <= SPRAYGPU_MAX_PRIMS, SPRAYGPU_ERR_LIMIT, "commit: %zu faces exceed the 23-bit primID field", nf);
  const int n = (int)nf;
  const int npairs = 3 * n;

  cudaEvent_t ev0, ev1;
  SG_CUDA(cudaEventCreate(&ev0));
  SG_CUDA(cudaEventCreate(&ev1));
  SG_CUDA(cudaEventRecord(ev0, st));

  // ---- persistent outputs
  int grc;
  if (sc->has_xf) {
    if ((grc = grow_buffer(&sc->d_verts_xf, &sc->cap_verts_xf, sizeof(float) * 3 * nv))) return grc;
    sc->d_verts = sc->d_verts_xf;
  } else sc->d_verts = sc->d_verts_obj;
  if ((grc = grow_buffer(&sc->d_normals, &sc->cap_normals, sizeof(float) * 3 * nv))) return grc;
  if ((grc = grow_buffer(&sc->d_tris, &sc->cap_tris, sizeof(float4) * 3 * nf))) return grc;
  if ((grc = grow_buffer(&sc->d_attr, &sc->cap_attr, sizeof(float4) * 3 * nf))) return grc;

  // ---- scratch carve (two passes: size, then pointers)
  const int nblocks_max = (npairs + RS_TILE - 1) / RS_TILE;
  uint32_t *kA, *kB, *vA, *vB, *hist, *vstart, *vend, *bounds;
  unsigned int* counters;
  float *tlo, *thi, *llo, *lhi, *blo, *bhi;
  int *bl, *br, *bfirst, *blast, *parent, *flags, *queue;
  uint4* tmp_nodes;
  auto carve = [&](Carver& c) {
    kA = c.take<uint32_t>(npairs); kB = c.take<uint32_t>(npairs);
    vA = c.take<uint32_t>(npairs); vB = c.take<uint32_t>(npairs);
    hist = c.take<uint32_t>((size_t)256 * nblocks_max + 1024);
    vstart = c.take<uint32_t>(nv); vend = c.take<uint32_t>(nv);
    bounds = c.take<uint32_t>(12); counters = c.take<unsigned int>(4);
    tlo = c.take<float>(3 * nf); thi = c.take<float>(3 * nf);
    llo = c.take<float>(3 * nf); lhi = c.take<float>(3 * nf);
    blo = c.take<float>(3 * nf); bhi = c.take<float>(3 * nf);
    bl = c.take<int>(nf); br = c.take<int>(nf); bfirst = c.take<int>(nf); blast = c.take<int>(nf);
    parent = c.take<int>(2 * nf); flags = c.take<int>(nf); queue = c.take<int>(nf);
    tmp_nodes = c.take<uint4>(5 * nf);
  };
  Carver sizer{nullptr};
  carve(sizer);
  int rc = ensure_scratch(dev->scratch, sizer.off + 256);
  if (rc) return rc;
  Carver carver{(char*)dev->scratch.ptr};
  carve(carver);

  const int T = 256;
  // ---- transform
  if (sc->has_xf) {
    Mat4 M; memcpy(M.m, sc->xf, sizeof(M.m));
    transform_kernel<<<grid_for(nv, T), T, 0, st>>>(sc->d_verts_obj, sc->d_verts, nv, M);
    dev->launches++;
  }
  // ---- normals (deterministic order)
  pairs_kernel<<<grid_for(npairs, T), T, 0, st>>>(sc->d_faces, kA, vA, npairs);
  dev->launches++;
  int vbits = 1; while (((size_t)1 << vbits) < nv) ++vbits;
  rc = radix_sort(dev, &kA, &vA, &kB, &vB, hist, npairs, vbits);
  if (rc) return rc;
  SG_CUDA(cudaMemsetAsync(vstart, 0, sizeof(uint32_t) * nv, st));
  SG_CUDA(cudaMemsetAsync(vend, 0, sizeof(uint32_t) * nv, st));
  segments_kernel<<<grid_for(npairs, T), T, 0, st>>>(kA, npairs, vstart, vend);
  normals_kernel<<<grid_for(nv, T), T, 0, st>>>(sc->d_verts, sc->d_faces, vA, vstart, vend, sc->d_normals, nv);
  attr_kernel<<<grid_for(nf, T), T, 0, st>>>(sc->d_faces, sc->d_normals, sc->d_rgb, n, sc->d_attr);
  dev->launches += 3;

  // ---- triangle boxes, Morton keys, sort
  {
    uint32_t hb[12];
    for (int a = 0; a < 3; ++a) {
      hb[a] = hb[6 + a] = 0xFFFFFFFFu;  // min slots
      hb[3 + a] = hb[9 + a] = 0u;       // max slots
    }
    SG_CUDA(cudaMemcpyAsync(bounds, hb, sizeof(hb), cudaMemcpyHostToDevice, st));
    // hb is on the stack: the copy from pageable memory is staged before the call returns
  }
  tri_bounds_kernel<<<std::min(grid_for(nf, 256), dev->sm_count * 8), 256, 0, st>>>(sc->d_verts, sc->d_faces, n, tlo, thi, bounds);
  dev->launches++;
  if (n == 1) {
    if ((grc = grow_buffer(&sc->d_nodes, &sc->cap_nodes, sizeof(uint4) * 5))) return grc;
    single_tri_kernel<<<1, 1, 0, st>>>(sc->d_verts, sc->d_faces, tlo, thi, sc->d_nodes, sc->d_tris);
    dev->launches++;
    sc->n_nodes = 1;
    sc->bvh_depth = 1;
  } else {
    morton_kernel<<<grid_for(nf, T), T, 0, st>>>(tlo, thi, n, bounds, kA, vA);
    dev->launches++;
    rc = radix_sort(dev, &kA, &vA, &kB, &vB, hist, n, 30);
    if (rc) return rc;
    leaf_boxes_kernel<<<grid_for(nf, T), T, 0, st>>>(vA, tlo, thi, n, llo, lhi);
    lbvh_kernel<<<grid_for(nf, T), T, 0, st>>>(kA, n, bl, br, bfirst, blast, parent);
    SG_CUDA(cudaMemsetAsync(flags, 0, sizeof(int) * nf, st));
    refit_kernel<<<grid_for(nf, T), T, 0, st>>>(parent, bl, br, llo, lhi, blo, bhi, flags, n);
    dev->launches += 3;
    // ---- collapse, level by level.  queue[j] = binary node of 8-wide node j; counters[0] = nodes
    // allocated so far, counters[1] = triangles emitted so far.
    {
      unsigned int hc[4] = {1u, 0u, 0u, 0u};
      int root = 0;
      SG_CUDA(cudaMemcpyAsync(counters, hc, sizeof(hc), cudaMemcpyHostToDevice, st));
      SG_CUDA(cudaMemcpyAsync(queue, &root, sizeof(int), cudaMemcpyHostToDevice, st));
      int begin = 0, end = 1, levels = 0;
      while (begin < end) {
        ++levels;
        collapse_kernel<<<grid_for(end - begin, 128), 128, 0, st>>>(begin, end, queue, counters, bl, br, bfirst,
                                                                    blast, blo, bhi, llo, lhi, vA, sc->d_verts,
                                                                    sc->d_faces, tmp_nodes, sc->d_tris);
        dev->launches++;
        unsigned int got[2];
        SG_CUDA(cudaMemcpyAsync(got, counters, sizeof(got), cudaMemcpyDeviceToHost, st));
        SG_CUDA(cudaStreamSynchronize(st));
        begin = end;
        end = (int)got[0];
        SG_CHECK(end <= n, SPRAYGPU_ERR_STATE, "collapse: node count %d exceeds bound %d", end, n);
      }
      sc->n_nodes = (uint32_t)end;
      sc->bvh_depth = (uint32_t)levels;
      // the traversal stack holds one entry per BVH8 level (core.cuh, SPRAY_STACK_SIZE): refuse deeper trees here
      // instead of dropping subtrees at trace time
      SG_CHECK(levels <= SPRAY_STACK_SIZE, SPRAYGPU_ERR_LIMIT, "commit: BVH8 depth %d exceeds the traversal stack (%d)", levels,
               SPRAY_STACK_SIZE);
    }
    if ((grc = grow_buffer(&sc->d_nodes, &sc->cap_nodes, sizeof(uint4) * 5 * (size_t)sc->n_nodes))) return grc;
    SG_CUDA(cudaMemcpyAsync(sc->d_nodes, tmp_nodes, sizeof(uint4) * 5 * (size_t)sc->n_nodes,
                            cudaMemcpyDeviceToDevice, st));
  }
  // ---- geometry bounds back to the host
  {
    uint32_t hb[12];
    SG_CUDA(cudaMemcpyAsync(hb, bounds, sizeof(hb), cudaMemcpyDeviceToHost, st));
    SG_CUDA(cudaEventRecord(ev1, st));
    SG_CUDA(cudaStreamSynchronize(st));
    for (int a = 0; a < 6; ++a) {
      uint32_t s = hb[6 + a];
      uint32_t u = (s & 0x80000000u) ? (s & 0x7fffffffu) : ~s;
      memcpy(&sc->bounds[a], &u, 4);
    }
  }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ev0, ev1);
  sc->build_ms = ms;
  cudaEventDestroy(ev0);
  cudaEventDestroy(ev1);
  SG_CUDA(cudaGetLastError());
  sc->committed = true;
  return SPRAYGPU_OK;
}

}  // namespace spraygpu
