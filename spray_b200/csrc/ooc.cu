// ooc.cu — the out-of-core blocking tile: spraygpu_trace_tile_ooc.
//
// Reference: ooc::PContext::isectPrims (src/ooc/ooc_pcontext.h:92-181) — per round: merge the per-domain ray
// counts, DomainStats::schedule() (src/ooc/ooc_domain_stats.cc:60-111), then for each scheduled domain
// `scene->load(id)` inside `omp single` (LruCache miss -> PLY parse + rtcCommit, :145-154) and drain that
// domain's radiance and shadow queues (TContext::procFilterQs).  Shadow rays of depth d and radiance rays of
// depth d+1 share a round (the speculation of ooc_tcontext.inl:52-135).
//
// Here the fused persistent kernel (trace.cu) needs every entered domain resident; when the device cache holds
// fewer domains than the tile enters, the tile runs as a WAVEFRONT swept over domain groups instead:
//   sweep s (s = 0 .. bounces):   radiance wave s (closest hit)  +  shadow rays emitted by wave s-1 (any hit)
//     1. count the rays entering each domain (exact slab test, aabb.h:139-169)            ooc_count_kernel
//     2. schedule: domains with rays, most rays first; cut into groups of <= max_resident
//     3. per group: load(user, id) for its domains (host callback = Scene::load: LRU hit, or PLY parse + commit
//        + bind); mask the table to the group; trace both ray sets against it, MERGING into the per-ray state:
//        64-bit atomicMin on the packed keys, OR on the occlusion bytes     intersect_/ooc_shadow_kernel
//        and capture colour + shading normal of every current winner whose domain is resident right now
//        (TriMeshBuffer::updateIntersection while the data is there)                      ooc_capture_kernel
//     4. retire the shadow rays (HdrImage::add where unoccluded)                           ooc_accumulate_kernel
//     5. shade wave s from the final keys + captured attributes (no domain data needed): all light-sample
//        shadow rays and the bounce ray -> wave s+1                                        ooc_shade_kernel
// so a frame costs (bounces + 1) sweeps over the entered domains, the same count as the reference's rounds.
// The result is independent of the grouping: closest hit = lexicographic min of (t, domain, prim).
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "internal.cuh"
#include "traverse.cuh"
#include "shade.cuh"

namespace spraygpu {

struct OocParams {
  const DomainEntry* table;
  int nd;
  spraygpu_shade shade;
  spraygpu_tile tile;
  int image_w, spp;
  int depth;     // depth of the wave being shaded
  int npasses;   // light-sample passes per diffuse hit
  size_t n;
};

__device__ __forceinline__ uint32_t ooc_pixid(const OocParams& P, size_t i) {
  const uint32_t pix = (uint32_t)(i / (size_t)P.spp);
  const uint32_t x0 = pix % (uint32_t)P.tile.w, y0 = pix / (uint32_t)P.tile.w;
  return (uint32_t)((P.tile.y + (int)y0) * P.image_w + P.tile.x + (int)x0);
}

__device__ __forceinline__ bool is_zero(vec3 d) { return d.x == 0.0f && d.y == 0.0f && d.z == 0.0f; }

__global__ void __launch_bounds__(256) ooc_fill_kernel(float* __restrict__ w, size_t n, float v) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) w[i] = v;
}

// step 1: rays per domain (DomainStats::increment, ooc_domain_stats.h).  Radiance rays of the current wave
// (when rad != 0) and the pending shadow rays (np passes; same origins) are counted together.
__global__ void __launch_bounds__(256)
    ooc_count_kernel(const DomainEntry* __restrict__ table, int nd, const spraygpu_ray* __restrict__ rays, size_t n, int rad,
                     const float* __restrict__ sdir, int np, unsigned int* __restrict__ counts) {
  extern __shared__ float smem[];
  float* boxes = smem;
  unsigned int* hist = (unsigned int*)(smem + 6 * nd);
  for (int i = threadIdx.x; i < nd; i += blockDim.x) {
#pragma unroll
    for (int k = 0; k < 6; ++k) boxes[6 * i + k] = table[i].box[k];
    hist[i] = 0u;
  }
  __syncthreads();
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) {
    const spraygpu_ray ry = load_ray(rays + (i));
    const vec3 o = mk3(ry.ox, ry.oy, ry.oz);
    for (int p = rad ? -1 : 0; p < np; ++p) {
      const vec3 d = p < 0 ? mk3(ry.dx, ry.dy, ry.dz)
                           : mk3(sdir[3 * ((size_t)p * n + i)], sdir[3 * ((size_t)p * n + i) + 1], sdir[3 * ((size_t)p * n + i) + 2]);
      if (is_zero(d)) continue;
      for (int k = 0; k < nd; ++k) {
        float tmin;
        if (domain_slab(boxes + 6 * k, o, d, SPRAY_RAY_EPS, INFINITY, &tmin)) atomicAdd(hist + k, 1u);
      }
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < nd; k += blockDim.x)
    if (hist[k]) atomicAdd(counts + k, hist[k]);
}

// step 3b: TriMeshBuffer::updateIntersection for every current winner whose domain is resident in this group.
// (u, v) are re-derived from (ray, prim) with the S1 arithmetic on the same vertex values (bit-identical to the
// traversal's), so only the key travels between groups.
__global__ void __launch_bounds__(128)
    ooc_capture_kernel(const DomainEntry* __restrict__ table, const spraygpu_ray* __restrict__ rays,
                       const uint64_t* __restrict__ keys, size_t n, uint32_t* __restrict__ color, float* __restrict__ ns3) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t key = keys[i];
  if (key == SPRAY_KEY_MISS) return;
  const DomainEntry* e = table + key_domain(key);
  if (!e->resident) return;
  const spraygpu_ray ry = load_ray(rays + (i));
  const uint32_t prim = key_prim(key);
  const uint32_t i0 = e->faces[3 * (size_t)prim], i1 = e->faces[3 * (size_t)prim + 1], i2 = e->faces[3 * (size_t)prim + 2];
  const vec3 v0 = mk3(e->verts[3 * (size_t)i0], e->verts[3 * (size_t)i0 + 1], e->verts[3 * (size_t)i0 + 2]);
  const vec3 v1 = mk3(e->verts[3 * (size_t)i1], e->verts[3 * (size_t)i1 + 1], e->verts[3 * (size_t)i1 + 2]);
  const vec3 v2 = mk3(e->verts[3 * (size_t)i2], e->verts[3 * (size_t)i2 + 1], e->verts[3 * (size_t)i2 + 2]);
  TriHit th; th.t = th.u = th.v = 0.f;
  tri_test(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), v0, gsub(v0, v1), gsub(v2, v0), ry.tmin, &th);
  uint32_t col; vec3 Ns;
  fetch_attributes(e, prim, th.u, th.v, &col, &Ns);
  color[i] = col;
  ns3[3 * i] = Ns.x; ns3[3 * i + 1] = Ns.y; ns3[3 * i + 2] = Ns.z;
}

// step 3c: pending shadow rays (pass-major: ray j = pass * n + sample) against the group's domains; occ |= hit
__global__ void __launch_bounds__(128)
    ooc_shadow_kernel(const DomainEntry* __restrict__ table, int nd, const TlasNode* __restrict__ tlas, int tlas_n,
                      const spraygpu_ray* __restrict__ rays, const float* __restrict__ sdir, size_t n, size_t total,
                      uint8_t* __restrict__ occ) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(table, nd, tlas, tlas_n, smem);
  const size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool valid = j < total && !occ[j];
  vec3 o = mk3(0, 0, 0), d = mk3(0, 0, 0);
  if (valid) {
    d = mk3(sdir[3 * j], sdir[3 * j + 1], sdir[3 * j + 2]);
    const spraygpu_ray ry = load_ray(rays + (j % n));   // the next wave's origin = this shadow ray's hit position
    o = mk3(ry.ox, ry.oy, ry.oz);
  }
  const bool live = valid && !is_zero(d);
  SceneHit hs;
  const bool oc = scene_trace_warp<false>(table, D, live, o, d, true, hs, nullptr);
  if (live && oc) occ[j] = 1;
}

// step 4: TContext::retire -> HdrImage::add (ooc_tcontext.inl:123-135)
__global__ void __launch_bounds__(256)
    ooc_accumulate_kernel(OocParams P, const float* __restrict__ slr, const uint8_t* __restrict__ occ, size_t total,
                          float* __restrict__ fb) {
  const size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (j >= total || occ[j]) return;
  const vec3 Lr = mk3(slr[3 * j], slr[3 * j + 1], slr[3 * j + 2]);
  if (!has_positive(Lr)) return;
  fb_add(fb, ooc_pixid(P, j % P.n), (float)(1.0 / (double)P.spp), Lr);
}

// step 5: shade wave `depth` from the final keys and the captured attributes (ShaderPt / ShaderAo,
// ooc_shader_pt.h:94-227, ooc_shader_ao.h:92-197).  Writes every light-sample shadow ray (pass-major) and the
// bounce record; samples that missed or terminated get zeros.
__global__ void __launch_bounds__(128)
    ooc_shade_kernel(OocParams P, const spraygpu_ray* __restrict__ rays, const uint64_t* __restrict__ keys,
                     const uint32_t* __restrict__ color, const float* __restrict__ ns3, const float* __restrict__ w,
                     float* __restrict__ sdir, float* __restrict__ slr, float* __restrict__ bnext,
                     unsigned long long* __restrict__ counters) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  uint32_t n_hit = 0, n_shadow = 0, n_rad = 0;
  if (i < P.n) {
    vec3 bd = mk3(0, 0, 0), bw = mk3(0, 0, 0);
    const uint64_t key = keys[i];
    const spraygpu_ray ry = load_ray(rays + (i));
    const vec3 ro = mk3(ry.ox, ry.oy, ry.oz), rd = mk3(ry.dx, ry.dy, ry.dz);
    const bool hit = key != SPRAY_KEY_MISS && !is_zero(rd);
    HitCtx c;
    const bool pt = P.shade.shader == SPRAYGPU_SHADER_PT;
    const int nl = pt ? P.shade.nlights : 1;
    const uint32_t samid = (uint32_t)i, pixid = ooc_pixid(P, i);
    const int next_depth = P.depth + 1;
    vec3 Lin = mk3(0, 0, 0);
    if (hit) {
      ++n_hit;
      const DomainEntry* e = P.table + key_domain(key);   // material lives in the table, resident or not
      const vec3 normal = mk3(ns3[3 * i], ns3[3 * i + 1], ns3[3 * i + 2]);
      c.pos = hit_position(ro, rd, key_t(key));
      c.kd = unpack_color(color[i]);
      c.wo = gneg(rd);
      c.cos_theta_i = gdot(c.wo, normal);
      c.entering = c.cos_theta_i > 0.0f;
      c.nff = gnormalize(c.entering ? normal : gneg(normal));
      c.mtl = e->mtl_type; c.eta0 = e->mtl[0]; c.eta1 = e->mtl[1];
      Lin = mk3(w[3 * i], w[3 * i + 1], w[3 * i + 2]);
      if (next_depth < P.shade.bounces) {
        bool em1, em2; vec3 d1 = mk3(0, 0, 1), w1 = mk3(0, 0, 0), d2 = mk3(0, 0, 1), w2 = mk3(0, 0, 0);
        shade_secondary<true>(c, Lin, samid, next_depth, &em1, &d1, &w1, &em2, &d2, &w2);
        // glass domains are rejected at launch, so at most one of the two is live
        if (em1) { bd = d1; bw = w1; ++n_rad; }
        else if (em2) { bd = d2; bw = w2; ++n_rad; }
      }
    }
    uint32_t ls = rng_init1(samid * (uint32_t)next_depth);   // ShaderPt light_sampler (ooc_shader_pt.h:129-130)
    int l = 0, k = 0;
    bool more = hit && c.mtl == SPRAYGPU_MTL_DIFFUSE && nl > 0;
    for (int p = 0; p < P.npasses; ++p) {
      vec3 sd = mk3(0, 0, 0), sl = mk3(0, 0, 0);
      if (more) {
        vec3 wi, Lr;
        if (shade_light_pass<true>(P.shade, pt, nl, c, Lin, pixid, l, k, ls, &wi, &Lr, &more)) { sd = wi; sl = Lr; ++n_shadow; }
      }
      const size_t j = (size_t)p * P.n + i;
      sdir[3 * j] = sd.x; sdir[3 * j + 1] = sd.y; sdir[3 * j + 2] = sd.z;
      slr[3 * j] = sl.x; slr[3 * j + 1] = sl.y; slr[3 * j + 2] = sl.z;
    }
    bnext[6 * i] = bd.x; bnext[6 * i + 1] = bd.y; bnext[6 * i + 2] = bd.z;
    bnext[6 * i + 3] = bw.x; bnext[6 * i + 4] = bw.y; bnext[6 * i + 5] = bw.z;
  }
  unsigned long long v[3] = {n_rad, n_shadow, n_hit};
  const int slots[3] = {2, 3, 7};
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    unsigned long long x = v[q];
#pragma unroll
    for (int off = 16; off; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
    if ((threadIdx.x & 31) == 0 && x) atomicAdd(counters + slots[q], x);
  }
}

// next wave: origin = hit position (also the origin of the pending shadow rays), direction = bounce (zero when
// the sample missed or terminated)
__global__ void __launch_bounds__(256)
    ooc_next_wave_kernel(spraygpu_ray* __restrict__ rays, const uint64_t* __restrict__ keys, const float* __restrict__ bnext,
                         float* __restrict__ w, size_t n) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  spraygpu_ray ry = load_ray(rays + (i));
  const vec3 rd = mk3(ry.dx, ry.dy, ry.dz);
  const uint64_t key = keys[i];
  vec3 o = mk3(0, 0, 0), d = mk3(0, 0, 0);
  if (key != SPRAY_KEY_MISS && !is_zero(rd)) {
    o = hit_position(mk3(ry.ox, ry.oy, ry.oz), rd, key_t(key));
    d = mk3(bnext[6 * i], bnext[6 * i + 1], bnext[6 * i + 2]);
  }
  ry.ox = o.x; ry.oy = o.y; ry.oz = o.z; ry.tmin = SPRAY_RAY_EPS;
  ry.dx = d.x; ry.dy = d.y; ry.dz = d.z; ry.tmax = INFINITY;
  store_ray(rays + i, ry);
  w[3 * i] = bnext[6 * i + 3]; w[3 * i + 1] = bnext[6 * i + 4]; w[3 * i + 2] = bnext[6 * i + 5];
}

}  // namespace spraygpu

using namespace spraygpu;

static inline int ooc_grid(size_t n, int t) { return (int)((n + t - 1) / t); }
// SPRAY_OOC_PREFETCH=0: the serial schedule of round 1 (A/B knob)
static inline bool ooc_prefetch_disabled() { const char* e = getenv("SPRAY_OOC_PREFETCH"); return e && *e == '0'; }

extern "C" int spraygpu_trace_tile_ooc(spraygpu_device dev_h, const spraygpu_camera* cam, const spraygpu_shade* shade,
                                       spraygpu_tile tile, float* fb, int max_resident, spraygpu_load_fn load, void* user,
                                       spraygpu_stats* stats) {
  SG_CHECK(dev_h && cam && shade && fb && load, SPRAYGPU_ERR_ARG, "trace_tile_ooc: NULL argument");
  SG_CHECK(max_resident >= 1, SPRAYGPU_ERR_ARG, "trace_tile_ooc: max_resident must be >= 1");
  Device* dev = (Device*)dev_h;
  SG_CHECK(shade->pixel_samples >= 1 && shade->bounces >= 1, SPRAYGPU_ERR_ARG, "trace_tile_ooc: pixel_samples and bounces must be >= 1");
  SG_CHECK(shade->nlights >= 0 && shade->nlights <= SPRAYGPU_MAX_LIGHTS, SPRAYGPU_ERR_LIMIT, "trace_tile_ooc: %d lights", shade->nlights);
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_AO || shade->nlights > 0, SPRAYGPU_ERR_ARG, "trace_tile_ooc: the PT shader needs a light");
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_PT || shade->ao_samples > 0, SPRAYGPU_ERR_ARG, "trace_tile_ooc: ao_samples must be > 0");
  SG_CHECK(tile.x >= 0 && tile.y >= 0 && tile.w >= 0 && tile.h >= 0 && tile.x + tile.w <= cam->image_w &&
               tile.y + tile.h <= cam->image_h, SPRAYGPU_ERR_ARG, "trace_tile_ooc: tile outside the image");
  SG_CHECK(dev->ndomains > 0, SPRAYGPU_ERR_STATE, "trace_tile_ooc: no domain table (spraygpu_domains_set)");
  SG_CUDA(cudaSetDevice(dev->cuda_device));
  {
    cudaPointerAttributes a;
    const bool devptr = cudaPointerGetAttributes(&a, fb) == cudaSuccess &&
                        (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    SG_CHECK(devptr, SPRAYGPU_ERR_ARG, "trace_tile_ooc: fb must be a device pointer");
  }
  for (int i = 0; i < dev->ndomains; ++i)
    SG_CHECK(dev->h_domains[i].mtl_type != SPRAYGPU_MTL_GLASS, SPRAYGPU_ERR_LIMIT,
             "trace_tile_ooc: glass domains (two secondary rays per hit) are not supported on the wavefront path; "
             "the reference CHECK-fails on them too (ooc_shader_pt.h:199-201)");
  const int nd = dev->ndomains;
  const size_t n = (size_t)tile.w * tile.h * shade->pixel_samples;
  cudaStream_t st = dev->stream;
  SG_CUDA(cudaMemsetAsync(dev->d_counters, 0, sizeof(unsigned long long) * 8, st));
  uint64_t scheduled_pairs = 0;
  int rc = SPRAYGPU_OK;
  if (n) {
    int npasses = 0;
    if (shade->shader == SPRAYGPU_SHADER_PT)
      for (int l = 0; l < shade->nlights; ++l) npasses += shade->lights[l].type == SPRAYGPU_LIGHT_DIFFUSE ? shade->ao_samples : 1;
    else npasses = shade->ao_samples;
    const size_t total = n * (size_t)npasses;

    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = (off + bytes + 255) & ~(size_t)255; return o; };
    const size_t o_rays = take(n * sizeof(spraygpu_ray)), o_keys = take(n * 8), o_w = take(n * 12), o_col = take(n * 4),
                 o_ns = take(n * 12), o_bnext = take(n * 24), o_sdir = take(total * 12), o_slr = take(total * 12),
                 o_occ = take(total), o_cnt = take(sizeof(unsigned int) * SPRAYGPU_MAX_DOMAINS);
    rc = ensure_scratch(dev->wave, off);
    if (rc) return rc;
    char* base = (char*)dev->wave.ptr;
    spraygpu_ray* rays = (spraygpu_ray*)(base + o_rays);
    uint64_t* keys = (uint64_t*)(base + o_keys);
    float* w = (float*)(base + o_w);
    uint32_t* col = (uint32_t*)(base + o_col);
    float* ns = (float*)(base + o_ns);
    float* bnext = (float*)(base + o_bnext);
    float* sdir = (float*)(base + o_sdir);
    float* slr = (float*)(base + o_slr);
    uint8_t* occ = (uint8_t*)(base + o_occ);
    unsigned int* d_cnt = (unsigned int*)(base + o_cnt);

    OocParams P;
    P.table = dev->d_domains; P.nd = nd; P.shade = *shade; P.tile = tile;
    P.image_w = cam->image_w; P.spp = shade->pixel_samples; P.n = n; P.npasses = npasses; P.depth = 0;

    // boxes and materials must be on the device before the first count (residency is refreshed per group)
    dev->active.clear();
    dev->table_dirty = true;
    rc = upload_domain_table(dev);
    if (rc) return rc;
    rc = launch_eye_rays(dev, cam, tile, shade->pixel_samples, shade->eye_seed, rays);
    if (rc) return rc;
    ooc_fill_kernel<<<ooc_grid(3 * n, 256), 256, 0, st>>>(w, 3 * n, 1.0f);
    dev->launches++;

    std::vector<std::pair<unsigned int, int>> order;
    for (int sweep = 0; sweep <= shade->bounces && rc == SPRAYGPU_OK; ++sweep) {
      const bool have_rad = sweep < shade->bounces;
      const bool have_sh = sweep > 0 && npasses > 0;
      if (!have_rad && !have_sh) break;
      if (have_rad) SG_CUDA(cudaMemsetAsync(keys, 0xFF, n * 8, st));
      if (have_sh) SG_CUDA(cudaMemsetAsync(occ, 0, total, st));
      // 1. rays per domain
      SG_CUDA(cudaMemsetAsync(d_cnt, 0, sizeof(unsigned int) * nd, st));
      ooc_count_kernel<<<ooc_grid(n, 256), 256, (size_t)nd * 28, st>>>(dev->d_domains, nd, rays, n, have_rad ? 1 : 0, sdir,
                                                                      have_sh ? npasses : 0, d_cnt);
      dev->launches++;
      unsigned int* h_cnt = (unsigned int*)dev->pinned;
      SG_CUDA(cudaMemcpyAsync(h_cnt, d_cnt, sizeof(unsigned int) * nd, cudaMemcpyDeviceToHost, st));
      SG_CUDA(cudaStreamSynchronize(st));
      // 2. schedule (DomainStats::schedule: most work first; ties by id so the order is deterministic)
      order.clear();
      for (int i = 0; i < nd; ++i)
        if (h_cnt[i]) { order.push_back(std::make_pair(h_cnt[i], i)); scheduled_pairs += h_cnt[i]; }
      std::sort(order.begin(), order.end(), [](const std::pair<unsigned int, int>& a, const std::pair<unsigned int, int>& b) {
        return a.first != b.first ? a.first > b.first : a.second < b.second;
      });
      // domains that are still resident from the previous sweep go first: their group costs no load at all (an
      // LRU cache walked in the same order every sweep would miss on every domain)
      std::stable_partition(order.begin(), order.end(), [&](const std::pair<unsigned int, int>& a) {
        return dev->h_domains[(size_t)a.second].resident != 0;
      });
      // 3. groups.  With room for two groups in the cache (max_resident >= 2) the NEXT group is loaded — mesh upload on
      // the upload stream, BVH build on the load stream — while the CURRENT one is traced on the device's stream; the
      // reference serialises Scene::load behind a barrier (ooc_pcontext.h:145-154).  An LRU cache of max_resident
      // blocks then evicts only domains of groups that were traced two steps ago: the load streams wait for that
      // trace's event before they overwrite a block.
      const bool prefetch = max_resident >= 2 && dev->load_stream && dev->upload_stream && !ooc_prefetch_disabled();
      const size_t gsz = prefetch ? (size_t)max_resident / 2 : (size_t)max_resident;
      const size_t ngroups = (order.size() + gsz - 1) / gsz;
      std::vector<cudaEvent_t> ev_trace(ngroups, nullptr);
      auto load_group = [&](size_t g) -> int {
        const size_t g0 = g * gsz, g1 = std::min(order.size(), g0 + gsz);
        if (prefetch) {
          // blocks evicted now were last read by the trace of group g - 2 at the latest
          if (g >= 2) {
            if (cudaStreamWaitEvent(dev->upload_stream, ev_trace[g - 2], 0) != cudaSuccess ||
                cudaStreamWaitEvent(dev->load_stream, ev_trace[g - 2], 0) != cudaSuccess) { set_error("trace_tile_ooc: cudaStreamWaitEvent failed"); return SPRAYGPU_ERR_CUDA; }
          }
          dev->loading_async = true;
        }
        int lrc2 = SPRAYGPU_OK;
        for (size_t g_ = g0; g_ < g1; ++g_) {
          const int id = order[g_].second;
          const int lrc = load(user, id);
          if (lrc) { set_error("trace_tile_ooc: load callback failed for domain %d (rc %d)", id, lrc); lrc2 = SPRAYGPU_ERR_STATE; break; }
        }
        dev->loading_async = false;
        return lrc2;
      };
      if (ngroups) rc = load_group(0);
      for (size_t g = 0; g < ngroups && rc == SPRAYGPU_OK; ++g) {
        const size_t g0 = g * gsz, g1 = std::min(order.size(), g0 + gsz);
        dev->active.assign((size_t)nd, 0);
        for (size_t g_ = g0; g_ < g1; ++g_) dev->active[(size_t)order[g_].second] = 1;
        for (size_t g_ = g0; g_ < g1; ++g_)
          if (!dev->h_domains[(size_t)order[g_].second].resident) {
            set_error("trace_tile_ooc: domain %d is not resident after its group was loaded (cache smaller than max_resident = %d?)",
                      order[g_].second, max_resident);
            rc = SPRAYGPU_ERR_STATE;
            break;
          }
        if (rc) break;
        dev->table_dirty = true;
        rc = upload_domain_table(dev);       // waits for the group's enqueued builds (scene_finalize), then uploads
        if (rc) break;
        if (have_rad) {
          rc = launch_intersect_domains(dev, rays, n, keys, nullptr, 1);
          if (rc) break;
          ooc_capture_kernel<<<ooc_grid(n, 128), 128, 0, st>>>(dev->d_domains, rays, keys, n, col, ns);
          dev->launches++;
        }
        if (have_sh) {
          SG_DYN_SMEM(ooc_shadow_kernel, dom_smem_bytes(nd, dev->tlas_n, 0));
          ooc_shadow_kernel<<<ooc_grid(total, 128), 128, dom_smem_bytes(nd, dev->tlas_n, 0), st>>>(
              dev->d_domains, nd, dev->d_tlas, dev->tlas_n, rays, sdir, n, total, occ);
          dev->launches++;
        }
        if (prefetch) {
          if (cudaEventCreateWithFlags(&ev_trace[g], cudaEventDisableTiming) != cudaSuccess || cudaEventRecord(ev_trace[g], st) != cudaSuccess) {
            set_error("trace_tile_ooc: cudaEventRecord failed"); rc = SPRAYGPU_ERR_CUDA; break;
          }
          // the cache must hold the group being traced AND the one being loaded: spraygpu_domain_bind refuses to
          // unbind (evict) a domain of the current group while the next one loads
          dev->tracing = dev->active;
          if (g + 1 < ngroups) rc = load_group(g + 1);       // overlaps the trace just enqueued
          dev->tracing.clear();
        } else if (g + 1 < ngroups) {
          rc = load_group(g + 1);                             // set_mesh waits for the device's stream: serial
        }
      }
      for (cudaEvent_t e : ev_trace) if (e) cudaEventDestroy(e);
      dev->active.clear();
      dev->table_dirty = true;
      if (rc) break;
      // 4. retire shadow rays of the previous wave
      if (have_sh) {
        ooc_accumulate_kernel<<<ooc_grid(total, 256), 256, 0, st>>>(P, slr, occ, total, fb);
        dev->launches++;
      }
      // 5. shade this wave, build the next one
      if (have_rad) {
        P.depth = sweep;
        ooc_shade_kernel<<<ooc_grid(n, 128), 128, 0, st>>>(P, rays, keys, col, ns, w, sdir, slr, bnext, dev->d_counters);
        ooc_next_wave_kernel<<<ooc_grid(n, 256), 256, 0, st>>>(rays, keys, bnext, w, n);
        dev->launches += 2;
      }
      SG_CUDA(cudaGetLastError());
    }
    dev->active.clear();
    dev->table_dirty = true;
    if (rc) return rc;
  }
  if (stats) {
    unsigned long long* h = (unsigned long long*)dev->pinned;
    SG_CUDA(cudaMemcpyAsync(h, dev->d_counters, sizeof(unsigned long long) * 8, cudaMemcpyDeviceToHost, st));
    SG_CUDA(cudaStreamSynchronize(st));
    stats->eye_rays = n; stats->radiance_rays = n + h[2]; stats->shadow_rays = h[3]; stats->domain_tests = scheduled_pairs;
    stats->nodes_visited = 0; stats->tris_tested = 0; stats->hits = h[7];
  }
  return SPRAYGPU_OK;
}
