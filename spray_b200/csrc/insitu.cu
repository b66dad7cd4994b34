// insitu.cu — the in-situ (domain-partitioned) blocking tile: spraygpu_trace_tile_insitu.
//
// Reference: insitu::{SingleThread,MultiThread}Tracer::trace — one super-step per ray depth
// (src/insitu/insitu_singlethread_tracer.inl:381-466): process rays in the local domains, composite the
// t-buffer with MPI_Allreduce(MIN) (insitu_vbuf.h:109-117) and the occlusion buffer with
// MPI_Allreduce(BOR) (:119-129), retire shadow rays, keep only secondaries whose parent t survived
// (:90-96).  Here a "wave" is the tile's samples at one depth, identical on every rank:
//   1. each rank traces the wave against ITS resident domains        (intersect_domains_kernel)
//   2. ncclAllReduce(uint64, min) on the packed (t | domain | prim) keys — carries the owner, so no
//      second round is needed to find who shades
//   3. the rank owning the winning domain shades: shadow direction + radiance, bounce ray  (insitu_shade_kernel)
//   4. all-reduce of the shadow directions (disjoint support -> integer sum of bit patterns)
//   5. every rank tests the shadow rays against its domains, ncclAllReduce(uint8, max)     (insitu_shadow_kernel)
//   6. the owner adds unoccluded radiance into ITS partial frame buffer                    (insitu_accumulate_kernel)
//   7. all-reduce of bounce direction + throughput -> next wave                            (insitu_next_wave_kernel)
#include "internal.cuh"
#include "traverse.cuh"
#include "shade.cuh"

namespace spraygpu {

struct Comm;
int comm_allreduce_u64_min(Comm* c, uint64_t* p, size_t n);
int comm_allreduce_u8_max(Comm* c, uint8_t* p, size_t n);
int comm_allreduce_bits_sum(Comm* c, uint32_t* p, size_t nwords);
int comm_allreduce_u64_sum(Comm* c, unsigned long long* p, size_t n);
int comm_rank(const Comm* c);
int comm_size(const Comm* c);
Device* comm_device(const Comm* c);

struct WaveParams {
  const DomainEntry* table;
  int nd;
  spraygpu_shade shade;
  spraygpu_tile tile;
  int image_w, spp;
  int depth;       // depth of the wave being shaded (eye rays: 0)
  int pass;        // light-sample pass index
  size_t n;
};

__device__ __forceinline__ uint32_t wave_pixid(const WaveParams& P, size_t i) {
  const uint32_t pix = (uint32_t)(i / (size_t)P.spp);
  const uint32_t x0 = pix % (uint32_t)P.tile.w, y0 = pix / (uint32_t)P.tile.w;
  return (uint32_t)((P.tile.y + (int)y0) * P.image_w + P.tile.x + (int)x0);
}

__global__ void __launch_bounds__(256) insitu_init_wave_kernel(float* __restrict__ w, size_t n) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < 3 * n) w[i] = 1.0f;
}

// step 3: the owner of the winning domain shades.  Everyone else writes zeros (the all-reduces rely on it).
__global__ void __launch_bounds__(128)
    insitu_shade_kernel(WaveParams P, const spraygpu_ray* __restrict__ rays, const uint64_t* __restrict__ keys,
                        const float* __restrict__ w, float* __restrict__ sdir, float* __restrict__ slr,
                        float* __restrict__ bnext /* 6n: dir, throughput; written when pass == 0 */,
                        unsigned long long* __restrict__ counters) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  uint32_t n_hit = 0, n_shadow = 0, n_rad = 0;
  if (i < P.n) {
    vec3 sd = mk3(0, 0, 0), sl = mk3(0, 0, 0), bd = mk3(0, 0, 0), bw = mk3(0, 0, 0);
    const uint64_t key = keys[i];
    if (key != SPRAY_KEY_MISS) {
      const DomainEntry* e = P.table + key_domain(key);
      if (e->resident) {
        const spraygpu_ray ry = load_ray(rays + (i));
        const vec3 ro = mk3(ry.ox, ry.oy, ry.oz), rd = mk3(ry.dx, ry.dy, ry.dz);
        const uint32_t prim = key_prim(key);
        // (u, v) re-derived from (ray, prim) with the S1 arithmetic on the same vertex values: bit-identical
        // to what the traversal computed, so only the key has to travel
        const uint32_t i0 = e->faces[3 * (size_t)prim], i1 = e->faces[3 * (size_t)prim + 1], i2 = e->faces[3 * (size_t)prim + 2];
        const vec3 v0 = mk3(e->verts[3 * (size_t)i0], e->verts[3 * (size_t)i0 + 1], e->verts[3 * (size_t)i0 + 2]);
        const vec3 v1 = mk3(e->verts[3 * (size_t)i1], e->verts[3 * (size_t)i1 + 1], e->verts[3 * (size_t)i1 + 2]);
        const vec3 v2 = mk3(e->verts[3 * (size_t)i2], e->verts[3 * (size_t)i2 + 1], e->verts[3 * (size_t)i2 + 2]);
        TriHit th; th.t = th.u = th.v = 0.f;
        tri_test(ro, rd, v0, gsub(v0, v1), gsub(v2, v0), ry.tmin, &th);
        HitCtx c;
        make_hit_ctx<true>(e, ro, rd, key_t(key), th.u, th.v, prim, c);
        const vec3 Lin = mk3(w[3 * i], w[3 * i + 1], w[3 * i + 2]);
        const uint32_t samid = (uint32_t)i, pixid = wave_pixid(P, i);
        const int next_depth = P.depth + 1;
        if (P.pass == 0) {
          ++n_hit;
          if (next_depth < P.shade.bounces) {
            bool em1, em2; vec3 d1 = mk3(0, 0, 1), w1 = mk3(0, 0, 0), d2 = mk3(0, 0, 1), w2 = mk3(0, 0, 0);
            shade_secondary<true>(c, Lin, samid, next_depth, &em1, &d1, &w1, &em2, &d2, &w2);
            // glass domains are rejected at launch, so at most one of the two is live
            if (em1) { bd = d1; bw = w1; ++n_rad; }
            else if (em2) { bd = d2; bw = w2; ++n_rad; }
          }
        }
        const bool pt = P.shade.shader == SPRAYGPU_SHADER_PT;
        const int nl = pt ? P.shade.nlights : 1;
        if (c.mtl == SPRAYGPU_MTL_DIFFUSE && nl > 0) {
          uint32_t ls = rng_init1(samid * (uint32_t)next_depth);
          int l = 0, k = 0; bool more = true;
          for (int p = 0; p <= P.pass && more; ++p) {
            vec3 wi, Lr;
            const bool emit = shade_light_pass<true>(P.shade, pt, nl, c, Lin, pixid, l, k, ls, &wi, &Lr, &more);
            if (p == P.pass && emit) { sd = wi; sl = Lr; ++n_shadow; }
          }
        }
      }
    }
    sdir[3 * i] = sd.x; sdir[3 * i + 1] = sd.y; sdir[3 * i + 2] = sd.z;
    slr[3 * i] = sl.x; slr[3 * i + 1] = sl.y; slr[3 * i + 2] = sl.z;
    if (P.pass == 0) {
      bnext[6 * i] = bd.x; bnext[6 * i + 1] = bd.y; bnext[6 * i + 2] = bd.z;
      bnext[6 * i + 3] = bw.x; bnext[6 * i + 4] = bw.y; bnext[6 * i + 5] = bw.z;
    }
  }
  // counters: [2] radiance rays, [3] shadow rays, [7] hits — warp-reduced
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

// step 5: shadow ray = (hit position from the reduced key, exchanged direction), any hit over the local domains
__global__ void __launch_bounds__(128)
    insitu_shadow_kernel(const DomainEntry* __restrict__ table, int nd, const TlasNode* __restrict__ tlas, int tlas_n,
                         const spraygpu_ray* __restrict__ rays,
                         const uint64_t* __restrict__ keys, const float* __restrict__ sdir, size_t n,
                         uint8_t* __restrict__ occ) {
  extern __shared__ float smem[];
  const DomSmem D = load_domain_smem(table, nd, tlas, tlas_n, smem);
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool valid = i < n;
  vec3 o = mk3(0, 0, 0), d = mk3(0, 0, 0);
  if (valid) {
    d = mk3(sdir[3 * i], sdir[3 * i + 1], sdir[3 * i + 2]);
    const uint64_t key = keys[i];
    if (key != SPRAY_KEY_MISS) {
      const spraygpu_ray ry = load_ray(rays + (i));
      o = hit_position(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), key_t(key));
    }
  }
  const bool live = valid && !(d.x == 0.0f && d.y == 0.0f && d.z == 0.0f);
  SceneHit hs;
  const bool oc = scene_trace_warp<false>(table, D, live, o, d, true, hs, nullptr);
  if (valid) occ[i] = oc ? 1 : 0;
}

// step 6: TContext::retireShadows -> HdrImage::add on the owner (slr is non-zero only there)
__global__ void __launch_bounds__(256)
    insitu_accumulate_kernel(WaveParams P, const float* __restrict__ slr, const uint8_t* __restrict__ occ,
                             float* __restrict__ fb) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= P.n || occ[i]) return;
  const vec3 Lr = mk3(slr[3 * i], slr[3 * i + 1], slr[3 * i + 2]);
  if (!has_positive(Lr)) return;
  fb_add(fb, wave_pixid(P, i), (float)(1.0 / (double)P.spp), Lr);
}

// step 7: next wave from the reduced bounce record; terminated samples get a zero direction
__global__ void __launch_bounds__(256)
    insitu_next_wave_kernel(spraygpu_ray* __restrict__ rays, const uint64_t* __restrict__ keys,
                            const float* __restrict__ bnext, float* __restrict__ w, size_t n) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  spraygpu_ray ry = load_ray(rays + (i));
  const vec3 d = mk3(bnext[6 * i], bnext[6 * i + 1], bnext[6 * i + 2]);
  const uint64_t key = keys[i];
  vec3 o = mk3(0, 0, 0);
  if (key != SPRAY_KEY_MISS) o = hit_position(mk3(ry.ox, ry.oy, ry.oz), mk3(ry.dx, ry.dy, ry.dz), key_t(key));
  ry.ox = o.x; ry.oy = o.y; ry.oz = o.z; ry.tmin = SPRAY_RAY_EPS;
  ry.dx = d.x; ry.dy = d.y; ry.dz = d.z; ry.tmax = INFINITY;
  store_ray(rays + i, ry);
  w[3 * i] = bnext[6 * i + 3]; w[3 * i + 1] = bnext[6 * i + 4]; w[3 * i + 2] = bnext[6 * i + 5];
}

}  // namespace spraygpu

using namespace spraygpu;

static inline int grid_for(size_t n, int t) { return (int)((n + t - 1) / t); }

extern "C" int spraygpu_trace_tile_insitu(spraygpu_device dev_h, spraygpu_comm comm_h, const spraygpu_camera* cam,
                                          const spraygpu_shade* shade, spraygpu_tile tile, float* fb,
                                          spraygpu_stats* stats) {
  SG_CHECK(dev_h && comm_h && cam && shade && fb, SPRAYGPU_ERR_ARG, "trace_tile_insitu: NULL argument");
  Comm* comm = (Comm*)comm_h;
  Device* dev = (Device*)dev_h;
  SG_CHECK(comm_device(comm) == dev, SPRAYGPU_ERR_ARG, "trace_tile_insitu: communicator belongs to another device");
  SG_CHECK(shade->pixel_samples >= 1 && shade->bounces >= 1, SPRAYGPU_ERR_ARG, "trace_tile_insitu: pixel_samples and bounces must be >= 1");
  SG_CHECK(shade->nlights >= 0 && shade->nlights <= SPRAYGPU_MAX_LIGHTS, SPRAYGPU_ERR_LIMIT, "trace_tile_insitu: %d lights", shade->nlights);
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_AO || shade->nlights > 0, SPRAYGPU_ERR_ARG, "trace_tile_insitu: the PT shader needs a light");
  SG_CHECK(shade->shader == SPRAYGPU_SHADER_PT || shade->ao_samples > 0, SPRAYGPU_ERR_ARG, "trace_tile_insitu: ao_samples must be > 0");
  SG_CHECK(tile.x >= 0 && tile.y >= 0 && tile.w >= 0 && tile.h >= 0 && tile.x + tile.w <= cam->image_w &&
               tile.y + tile.h <= cam->image_h, SPRAYGPU_ERR_ARG, "trace_tile_insitu: tile outside the image");
  SG_CUDA(cudaSetDevice(dev->cuda_device));
  {
    cudaPointerAttributes a;
    const bool devptr = cudaPointerGetAttributes(&a, fb) == cudaSuccess &&
                        (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    SG_CHECK(devptr, SPRAYGPU_ERR_ARG, "trace_tile_insitu: fb must be a device pointer");
  }
  for (int i = 0; i < dev->ndomains; ++i)
    SG_CHECK(dev->h_domains[i].mtl_type != SPRAYGPU_MTL_GLASS, SPRAYGPU_ERR_LIMIT,
             "trace_tile_insitu: glass domains (two secondary rays per hit) are not supported on the in-situ path; "
             "the reference CHECK-fails on them too (insitu_shader_pt.h)");
  int rc = upload_domain_table(dev);
  if (rc) return rc;
  const size_t n = (size_t)tile.w * tile.h * shade->pixel_samples;
  cudaStream_t st = dev->stream;
  SG_CUDA(cudaMemsetAsync(dev->d_counters, 0, sizeof(unsigned long long) * 8, st));
  if (n) {
    // wave buffers (device scratch, grown on demand)
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = (off + bytes + 255) & ~(size_t)255; return o; };
    const size_t o_rays = take(n * sizeof(spraygpu_ray)), o_keys = take(n * 8), o_w = take(n * 12), o_sdir = take(n * 12),
                 o_slr = take(n * 12), o_bnext = take(n * 24), o_occ = take(n);
    rc = ensure_scratch(dev->wave, off);
    if (rc) return rc;
    char* base = (char*)dev->wave.ptr;
    spraygpu_ray* rays = (spraygpu_ray*)(base + o_rays);
    uint64_t* keys = (uint64_t*)(base + o_keys);
    float* w = (float*)(base + o_w);
    float* sdir = (float*)(base + o_sdir);
    float* slr = (float*)(base + o_slr);
    float* bnext = (float*)(base + o_bnext);
    uint8_t* occ = (uint8_t*)(base + o_occ);

    WaveParams P;
    P.table = dev->d_domains; P.nd = dev->ndomains; P.shade = *shade; P.tile = tile;
    P.image_w = cam->image_w; P.spp = shade->pixel_samples; P.n = n;
    // light-sample passes per diffuse hit
    int npasses = 0;
    if (shade->shader == SPRAYGPU_SHADER_PT)
      for (int l = 0; l < shade->nlights; ++l) npasses += shade->lights[l].type == SPRAYGPU_LIGHT_DIFFUSE ? shade->ao_samples : 1;
    else npasses = shade->ao_samples;

    rc = launch_eye_rays(dev, cam, tile, shade->pixel_samples, SPRAYGPU_EYE_SEED_INSITU, rays);   // insitu_ray.h:161-165
    if (rc) return rc;
    insitu_init_wave_kernel<<<grid_for(3 * n, 256), 256, 0, st>>>(w, n);
    dev->launches++;
    const size_t dsmem = dom_smem_bytes(dev->ndomains, dev->tlas_n, 0);
    SG_DYN_SMEM(insitu_shadow_kernel, dsmem);
    for (int depth = 0; depth < shade->bounces; ++depth) {
      P.depth = depth;
      rc = launch_intersect_domains(dev, rays, n, keys, nullptr, 0);                       // 1
      if (rc) return rc;
      rc = comm_allreduce_u64_min(comm, keys, n);                                          // 2
      if (rc) return rc;
      for (int pass = 0; pass < (npasses > 0 ? npasses : 1); ++pass) {
        P.pass = pass;
        insitu_shade_kernel<<<grid_for(n, 128), 128, 0, st>>>(P, rays, keys, w, sdir, slr, bnext, dev->d_counters);   // 3
        dev->launches++;
        if (npasses == 0) break;
        rc = comm_allreduce_bits_sum(comm, (uint32_t*)sdir, 3 * n);                        // 4
        if (rc) return rc;
        insitu_shadow_kernel<<<grid_for(n, 128), 128, dsmem, st>>>(dev->d_domains, dev->ndomains, dev->d_tlas, dev->tlas_n, rays, keys, sdir, n, occ);  // 5
        dev->launches++;
        rc = comm_allreduce_u8_max(comm, occ, n);
        if (rc) return rc;
        insitu_accumulate_kernel<<<grid_for(n, 256), 256, 0, st>>>(P, slr, occ, fb);       // 6
        dev->launches++;
      }
      if (depth + 1 < shade->bounces) {
        rc = comm_allreduce_bits_sum(comm, (uint32_t*)bnext, 6 * n);                       // 7
        if (rc) return rc;
        insitu_next_wave_kernel<<<grid_for(n, 256), 256, 0, st>>>(rays, keys, bnext, w, n);
        dev->launches++;
      }
    }
    SG_CUDA(cudaGetLastError());
  }
  if (stats) {
    // owners counted hits, shadow and secondary rays; eye rays (n) are identical on every rank
    rc = comm_allreduce_u64_sum(comm, dev->d_counters, 8);
    if (rc) return rc;
    unsigned long long* h = (unsigned long long*)dev->pinned;
    SG_CUDA(cudaMemcpyAsync(h, dev->d_counters, sizeof(unsigned long long) * 8, cudaMemcpyDeviceToHost, st));
    SG_CUDA(cudaStreamSynchronize(st));
    stats->eye_rays = n; stats->radiance_rays = n + h[2]; stats->shadow_rays = h[3]; stats->domain_tests = 0;
    stats->nodes_visited = 0; stats->tris_tested = 0; stats->hits = h[7];
  }
  return SPRAYGPU_OK;
}
