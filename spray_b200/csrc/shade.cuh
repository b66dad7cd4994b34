// shade.cuh — shading of one radiance hit, shared by the fused persistent kernel (trace.cu) and the
// in-situ wavefront kernels (insitu.cu) so both produce bit-identical rays and radiance.
// Reference: ShaderPt / ShaderAo (src/ooc/ooc_shader_pt.h:94-227, src/ooc/ooc_shader_ao.h:92-197,
// src/insitu/insitu_shader_pt.h:83-230), TriMeshBuffer::updateIntersection (trimesh_buffer.cc:328-360).
#pragma once
#include "internal.cuh"

namespace spraygpu {

// TriMeshBuffer::updateIntersection (trimesh_buffer.cc:328-360) from the 48-byte attribute record
__device__ __forceinline__ void fetch_attributes(const DomainEntry* e, uint32_t prim, float u, float v,
                                                 uint32_t* color, vec3* Ns) {
  const float4* p = e->attr + 3 * (size_t)prim;
  const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  const float w = sub_rn(sub_rn(1.f, u), v);
  *color = interp_color(__float_as_uint(a.w), __float_as_uint(b.w), __float_as_uint(c.w), u, v, w);
  Ns->x = add_rn(add_rn(mul_rn(a.x, w), mul_rn(b.x, u)), mul_rn(c.x, v));
  Ns->y = add_rn(add_rn(mul_rn(a.y, w), mul_rn(b.y, u)), mul_rn(c.y, v));
  Ns->z = add_rn(add_rn(mul_rn(a.z, w), mul_rn(b.z, u)), mul_rn(c.z, v));
}

// hit position: org + dir * t, one rounding per op (ooc_shader_pt.h:104)
__device__ __forceinline__ vec3 hit_position(vec3 ro, vec3 rd, float t) {
  return mk3(add_rn(mul_rn(rd.x, t), ro.x), add_rn(mul_rn(rd.y, t), ro.y), add_rn(mul_rn(rd.z, t), ro.z));
}

struct HitCtx {
  vec3 pos, kd, nff, wo;
  float cos_theta_i, eta0, eta1;
  bool entering;
  int mtl;
};

template <bool GENERAL>
__device__ __forceinline__ void make_hit_ctx(const DomainEntry* e, vec3 ro, vec3 rd, float t, float u, float v,
                                             uint32_t prim, HitCtx& c) {
  uint32_t color; vec3 normal;
  fetch_attributes(e, prim, u, v, &color, &normal);
  c.pos = hit_position(ro, rd, t);
  c.kd = unpack_color(color);
  c.wo = gneg(rd);
  c.cos_theta_i = gdot(c.wo, normal);
  c.entering = c.cos_theta_i > 0.0f;
  c.nff = gnormalize(c.entering ? normal : gneg(normal));
  c.mtl = SPRAYGPU_MTL_DIFFUSE; c.eta0 = 1.f; c.eta1 = 1.f;
  if (GENERAL) { c.mtl = e->mtl_type; c.eta0 = e->mtl[0]; c.eta1 = e->mtl[1]; }
}

// One light-sample pass of a hit: light l, sample k of it (area lights / AO take ao_samples samples,
// point lights one).  Advances (l, k) and the light sampler state; returns whether a shadow ray is
// emitted (direction wi, radiance Lr to add if unoccluded).  `more` = further passes remain.
template <bool GENERAL>
__device__ __forceinline__ bool shade_light_pass(const spraygpu_shade& S, bool pt, int nl, const HitCtx& c, vec3 Lin,
                                                 uint32_t pixid, int& l, int& k, uint32_t& ls, vec3* wi_o, vec3* Lr_o,
                                                 bool* more) {
  const vec3 ks = mk3(S.ks[0], S.ks[1], S.ks[2]);
  const spraygpu_light& L = S.lights[pt ? l : 0];
  const bool area = GENERAL ? (pt ? (L.type == SPRAYGPU_LIGHT_DIFFUSE) : true) : false;
  const vec3 rad = mk3(L.radiance[0], L.radiance[1], L.radiance[2]);
  bool emit = false; vec3 wi = mk3(0, 0, 1), Lr = mk3(0, 0, 0);
  float costheta;
  if (GENERAL && area) {
    float pdf;
    if (pt) {
      const float u1 = rng_float(ls), u2 = rng_float(ls);
      cosine_hemisphere(u1, u2, c.nff, &wi, &pdf);
      if (pdf > 0.0f) {
        costheta = gclamp(gdot(c.nff, wi), 0.0f, 1.0f);
        Lr = gscale(gmul(Lin, blinn_phong(costheta, c.kd, ks, S.shininess, rad, wi, c.nff, c.wo)),
                    div_rn(1.0f, mul_rn(pdf, (float)S.ao_samples)));
        emit = true;
      }
    } else {
      const float ao_weight = div_rn(1.0f, (float)S.ao_samples);
      uint32_t as = rng_init1(pixid * (uint32_t)(k + 1));  // ooc_shader_ao.h:131
      const float u1 = rng_float(as), u2 = rng_float(as);
      cosine_hemisphere(u1, u2, c.nff, &wi, &pdf);
      costheta = gclamp(gdot(c.nff, wi), 0.0f, 1.0f);
      Lr = gscale(gmul(Lin, c.kd), div_rn(mul_rn(mul_rn(SPRAY_ONE_OVER_PI, costheta), ao_weight), pdf));
      emit = true;
    }
  } else {
    wi = gnormalize(gsub(mk3(L.position[0], L.position[1], L.position[2]), c.pos));
    costheta = gclamp(gdot(c.nff, wi), 0.0f, 1.0f);
    Lr = gscale(gmul(Lin, blinn_phong(costheta, c.kd, ks, S.shininess, rad, wi, c.nff, c.wo)), 1.0f);
    emit = true;
  }
  emit = emit && has_positive(Lr);
  ++k;
  if (k >= (area ? S.ao_samples : 1)) { k = 0; ++l; }
  *more = l < nl;
  *wi_o = wi; *Lr_o = Lr;
  return emit;
}

// Secondary rays of a hit (bounce): up to two (glass: reflection d1/w1 and transmission d2/w2).
// The reference CHECK-fails when glass yields both (ooc_shader_pt.h:199-201); the fused kernel traces both.
template <bool GENERAL>
__device__ __forceinline__ void shade_secondary(const HitCtx& c, vec3 Lin, uint32_t samid, int next_depth, bool* em1,
                                                vec3* d1, vec3* w1, bool* em2, vec3* d2, vec3* w2) {
  *em1 = false; *em2 = false;
  const vec3 won = gnormalize(c.wo);
  if (GENERAL && c.mtl != SPRAYGPU_MTL_DIFFUSE) {
    if (c.cos_theta_i != 0.0f) {
      float ci = gclamp(c.cos_theta_i, -1.0f, 1.0f);
      const float abs_cos = fabsf(ci);
      if (!c.entering) ci = abs_cos;
      bool has_reflect = false, has_trans = false; float fr = 0.f; vec3 wt = mk3(0, 0, 0);
      if (c.mtl == SPRAYGPU_MTL_MIRROR) { fr = 1.0f; has_reflect = true; }
      else {
        const float etaI = c.entering ? c.eta0 : c.eta1, etaT = c.entering ? c.eta1 : c.eta0;
        if (c.mtl == SPRAYGPU_MTL_GLASS) {
          fr = fr_dielectric(ci, etaI, etaT, won, c.nff, &wt);
          if (fr == 1.0f) has_reflect = true;
          else if (fr == 0.0f) has_trans = true;
          else { has_reflect = true; has_trans = true; }
        } else {
          if (refract_dir(ci, etaI, etaT, won, c.nff, &wt)) { has_trans = true; fr = 0.0f; }
          else { has_reflect = true; fr = 1.0f; }
        }
      }
      if (has_reflect) {
        const vec3 wr = gadd(gneg(won), gscale(c.nff, mul_rn(2.0f, gdot(won, c.nff))));
        *d1 = gnormalize(wr);
        *w1 = gscale(Lin, div_rn(fr, abs_cos));
        *em1 = has_positive(*w1);
      }
      if (has_trans) {
        *d2 = gnormalize(wt);
        *w2 = gscale(Lin, div_rn(sub_rn(1.0f, fr), abs_cos));
        *em2 = has_positive(*w2);
      }
    }
  } else {
    uint32_t rs = rng_init1(samid * (uint32_t)next_depth);
    const float u1 = rng_float(rs), u2 = rng_float(rs);
    float pdf;
    cosine_hemisphere(u1, u2, c.nff, d1, &pdf);
    const float costheta = gclamp(gdot(c.nff, *d1), 0.0f, 1.0f);
    const vec3 Lr = gscale(gscale(gmul(Lin, c.kd), SPRAY_ONE_OVER_PI), costheta);
    *w1 = mk3(div_rn(Lr.x, pdf), div_rn(Lr.y, pdf), div_rn(Lr.z, pdf));
    *em1 = has_positive(*w1);
  }
}

__device__ __forceinline__ void fb_add(float* fb, uint32_t pixid, float scale, vec3 Lr) {
  // TContext::retire -> HdrImage::add (ooc_tcontext.inl:123-135, image.h:99-109)
  float* px = fb + 4 * (size_t)pixid;
  atomicAdd(px + 0, mul_rn(scale, Lr.x));
  atomicAdd(px + 1, mul_rn(scale, Lr.y));
  atomicAdd(px + 2, mul_rn(scale, Lr.z));
}

// tile = the rectangle being rendered; (pw, px, py) place it inside the BLOCKING tile that numbers the samples
// (samid = index within the blocking tile, ooc_tracer.inl:84-172): pw = blocking tile width, (px, py) = offset of
// `tile` inside it.  A rank that renders only some rows of a blocking tile (multi-GPU image partition) still seeds
// every sample exactly as the one-rank frame does.
__device__ __forceinline__ void eye_sample(const CameraDev& cam, spraygpu_tile tile, int pw, int px, int py, int spp, int eye_seed,
                                           uint32_t x0, uint32_t y0, uint32_t sidx, uint32_t* samid, uint32_t* pixid, vec3* dir) {
  const uint32_t bufid = (uint32_t)spp * ((y0 + (uint32_t)py) * (uint32_t)pw + x0 + (uint32_t)px) + sidx;
  const int x = tile.x + (int)x0, y = tile.y + (int)y0;
  float fx = (float)x, fy = (float)y;
  if (spp > 1) {
    // ooc::Tracer::genMultiEyes seeds with bufid (ooc_tracer.inl:151-155); insitu::genMultiSampleEyeRays with
    // (pixid, s) (insitu_ray.h:161-165)
    uint32_t s = eye_seed == SPRAYGPU_EYE_SEED_INSITU ? rng_init2((uint32_t)(y * cam.image_w + x), sidx) : rng_init1(bufid);
    fx = add_rn((float)x, rng_float(s));
    fy = add_rn((float)y, rng_float(s));
  }
  *dir = camera_dir(cam, fx, fy);
  *samid = bufid;
  *pixid = (uint32_t)(y * cam.image_w + x);
}

}  // namespace spraygpu
