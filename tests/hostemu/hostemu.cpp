// hostemu.cpp — DEVELOPMENT/TEST harness: runs the __host__ __device__ building blocks of
// spray_b200/csrc/core.cuh serially on the CPU (LBVH -> 8-wide collapse -> traversal) so the logic can
// be checked against the oracle in the GPU-less build container.  It is NOT a product path and is
// never loaded by spray_b200/; tests/test_hostemu.py is its only user.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../spray_b200/csrc/core.cuh"

using namespace spraygpu;

struct EmuScene {
  std::vector<u4> nodes;        // 5 per node
  std::vector<float> tris;      // 12 floats per tri (leaf order)
  std::vector<float> verts;
  std::vector<uint32_t> faces;
  int max_sp = 0;
  int depth = 1;                // BVH8 levels (what spraygpu_scene_commit checks against SPRAY_STACK_SIZE)
};

struct NodeLoaderH {
  const u4* nodes;
  void operator()(uint32_t idx, u4& q0, u4& q1, u4& q2, u4& q3, u4& q4) const {
    const u4* p = nodes + 5 * (size_t)idx; q0 = p[0]; q1 = p[1]; q2 = p[2]; q3 = p[3]; q4 = p[4];
  }
};
struct TriLoaderH {
  const float* tris;
  void operator()(uint32_t idx, vec3& v0, vec3& e1, vec3& e2, uint32_t& prim) const {
    const float* p = tris + 12 * (size_t)idx;
    v0 = mk3(p[0], p[1], p[2]); memcpy(&prim, p + 3, 4);
    e1 = mk3(p[4], p[5], p[6]); e2 = mk3(p[8], p[9], p[10]);
  }
};

static void write_tri(const EmuScene& s, uint32_t prim, float* out) {
  const uint32_t* f = &s.faces[3 * (size_t)prim];
  vec3 v0 = mk3(s.verts[3 * f[0]], s.verts[3 * f[0] + 1], s.verts[3 * f[0] + 2]);
  vec3 v1 = mk3(s.verts[3 * f[1]], s.verts[3 * f[1] + 1], s.verts[3 * f[1] + 2]);
  vec3 v2 = mk3(s.verts[3 * f[2]], s.verts[3 * f[2] + 1], s.verts[3 * f[2] + 2]);
  vec3 e1 = gsub(v0, v1), e2 = gsub(v2, v0);
  out[0] = v0.x; out[1] = v0.y; out[2] = v0.z; memcpy(out + 3, &prim, 4);
  out[4] = e1.x; out[5] = e1.y; out[6] = e1.z; out[7] = 0;
  out[8] = e2.x; out[9] = e2.y; out[10] = e2.z; out[11] = 0;
}

// Binned-SAH top-down binary builder (16 bins on the widest centroid axis... all three axes tried), used only to
// MEASURE how far the shipped Morton LBVH is from a SAH-quality tree (tools/bvh_quality.py, EMU_BUILDER=sah).
struct SahCtx {
  const float* tlo; const float* thi;      // per-primitive boxes
  std::vector<uint32_t>* order;            // permutation being partitioned in place
  std::vector<int>*bl, *br, *bfirst, *blast;
  std::vector<float>*blo, *bhi;
  int next = 0;
};
static float half_area(const float* lo, const float* hi) {
  float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
  return dx * dy + dy * dz + dz * dx;
}
static int sah_build(SahCtx& c, int first, int last, int self) {   // returns child ref; self < 0 -> allocate
  if (first == last) return ~first;
  const int me = self >= 0 ? self : c.next++;
  std::vector<uint32_t>& ord = *c.order;
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = first; i <= last; ++i)
    for (int a = 0; a < 3; ++a) {
      float l = c.tlo[3 * (size_t)ord[i] + a], h = c.thi[3 * (size_t)ord[i] + a], m = 0.5f * (l + h);
      lo[a] = fminf(lo[a], l); hi[a] = fmaxf(hi[a], h); clo[a] = fminf(clo[a], m); chi[a] = fmaxf(chi[a], m);
    }
  for (int a = 0; a < 3; ++a) { (*c.blo)[3 * (size_t)me + a] = lo[a]; (*c.bhi)[3 * (size_t)me + a] = hi[a]; }
  (*c.bfirst)[me] = first; (*c.blast)[me] = last;
  const int NB = 16;
  float best = INFINITY; int best_axis = -1, best_bin = -1;
  for (int a = 0; a < 3; ++a) {
    float ext = chi[a] - clo[a];
    if (!(ext > 0.f)) continue;
    float bl_[NB][3], bh_[NB][3]; int bc[NB];
    for (int b = 0; b < NB; ++b) { bc[b] = 0; for (int k = 0; k < 3; ++k) { bl_[b][k] = INFINITY; bh_[b][k] = -INFINITY; } }
    for (int i = first; i <= last; ++i) {
      const float* l = c.tlo + 3 * (size_t)ord[i]; const float* h = c.thi + 3 * (size_t)ord[i];
      int b = (int)((0.5f * (l[a] + h[a]) - clo[a]) / ext * NB); b = b < 0 ? 0 : (b >= NB ? NB - 1 : b);
      bc[b]++;
      for (int k = 0; k < 3; ++k) { bl_[b][k] = fminf(bl_[b][k], l[k]); bh_[b][k] = fmaxf(bh_[b][k], h[k]); }
    }
    float ra[NB]; int rc[NB];
    float al[3] = {INFINITY, INFINITY, INFINITY}, ah[3] = {-INFINITY, -INFINITY, -INFINITY}; int cnt = 0;
    for (int b = NB - 1; b > 0; --b) {
      for (int k = 0; k < 3; ++k) { al[k] = fminf(al[k], bl_[b][k]); ah[k] = fmaxf(ah[k], bh_[b][k]); }
      cnt += bc[b]; ra[b] = cnt ? half_area(al, ah) : 0.f; rc[b] = cnt;
    }
    for (int k = 0; k < 3; ++k) { al[k] = INFINITY; ah[k] = -INFINITY; }
    cnt = 0;
    for (int b = 0; b < NB - 1; ++b) {
      for (int k = 0; k < 3; ++k) { al[k] = fminf(al[k], bl_[b][k]); ah[k] = fmaxf(ah[k], bh_[b][k]); }
      cnt += bc[b];
      if (cnt == 0 || rc[b + 1] == 0) continue;
      float cost = half_area(al, ah) * (float)cnt + ra[b + 1] * (float)rc[b + 1];
      if (cost < best) { best = cost; best_axis = a; best_bin = b; }
    }
  }
  int mid;
  if (best_axis < 0) {
    mid = (first + last) / 2;       // coincident centroids: split the range in the middle
  } else {
    const int a = best_axis; const float ext = chi[a] - clo[a];
    auto it = std::stable_partition(ord.begin() + first, ord.begin() + last + 1, [&](uint32_t p) {
      int b = (int)((0.5f * (c.tlo[3 * (size_t)p + a] + c.thi[3 * (size_t)p + a]) - clo[a]) / ext * NB);
      b = b < 0 ? 0 : (b >= NB ? NB - 1 : b);
      return b <= best_bin; });
    mid = (int)(it - ord.begin()) - 1;
  }
  int l = sah_build(c, first, mid, -1);
  int r = sah_build(c, mid + 1, last, -1);
  (*c.bl)[me] = l; (*c.br)[me] = r;
  return me;
}

extern "C" {

EmuScene* emu_build(const float* verts, size_t nv, const uint32_t* faces, size_t nf) {
  EmuScene* s = new EmuScene();
  s->verts.assign(verts, verts + 3 * nv);
  s->faces.assign(faces, faces + 3 * nf);
  const int n = (int)nf;
  std::vector<float> tlo(3 * nf), thi(3 * nf);
  float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (size_t i = 0; i < nf; ++i) {
    for (int a = 0; a < 3; ++a) { tlo[3 * i + a] = INFINITY; thi[3 * i + a] = -INFINITY; }
    for (int k = 0; k < 3; ++k)
      for (int a = 0; a < 3; ++a) {
        float x = verts[3 * (size_t)faces[3 * i + k] + a];
        tlo[3 * i + a] = fminf(tlo[3 * i + a], x); thi[3 * i + a] = fmaxf(thi[3 * i + a], x);
      }
    for (int a = 0; a < 3; ++a) {
      float c = 0.5f * (tlo[3 * i + a] + thi[3 * i + a]);
      clo[a] = fminf(clo[a], c); chi[a] = fmaxf(chi[a], c);
    }
  }
  s->tris.resize(12 * nf);
  if (n == 1) {
    CollapseOut co; co.n_inner = 0; co.n_tris = 1; co.n_leaf = 1; co.imask = 0;
    for (int k = 0; k < 8; ++k) co.meta[k] = 0;
    co.meta[0] = 1u << 5; co.tri_first[0] = 0; co.tri_count[0] = 1;
    for (int a = 0; a < 3; ++a) { co.lo[0][a] = tlo[a]; co.hi[0][a] = thi[a]; }
    s->nodes.resize(5);
    encode_node8(tlo.data(), thi.data(), co, 0, 0, s->nodes.data());
    write_tri(*s, 0, s->tris.data());
    return s;
  }
  std::vector<uint32_t> keys(nf), order(nf);
  for (size_t i = 0; i < nf; ++i) {
    float p[3];
    for (int a = 0; a < 3; ++a) {
      float c = 0.5f * (tlo[3 * i + a] + thi[3 * i + a]);
      float ext = chi[a] - clo[a];
      p[a] = ext > 0.f ? (c - clo[a]) / ext : 0.f;
    }
    keys[i] = morton30(p[0], p[1], p[2]);
  }
  std::iota(order.begin(), order.end(), 0u);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
  std::vector<uint32_t> skeys(nf);
  std::vector<float> llo(3 * nf), lhi(3 * nf);
  for (size_t i = 0; i < nf; ++i) {
    skeys[i] = keys[order[i]];
    for (int a = 0; a < 3; ++a) { llo[3 * i + a] = tlo[3 * (size_t)order[i] + a]; lhi[3 * i + a] = thi[3 * (size_t)order[i] + a]; }
  }
  std::vector<int> bl(n), br(n), bfirst(n), blast(n), parent(2 * n, -1);
  std::vector<float> blo(3 * nf, INFINITY), bhi(3 * nf, -INFINITY);
  const char* builder = getenv("EMU_BUILDER");
  const bool sah = builder && !strcmp(builder, "sah");
  if (sah) {
    SahCtx c{tlo.data(), thi.data(), &order, &bl, &br, &bfirst, &blast, &blo, &bhi, 1};
    std::iota(order.begin(), order.end(), 0u);
    sah_build(c, 0, n - 1, 0);
    for (size_t i = 0; i < nf; ++i)
      for (int a = 0; a < 3; ++a) { llo[3 * i + a] = tlo[3 * (size_t)order[i] + a]; lhi[3 * i + a] = thi[3 * (size_t)order[i] + a]; }
  }
  for (int i = 0; i < n - 1 && !sah; ++i) {
    lbvh_node(skeys.data(), n, i, &bl[i], &br[i], &bfirst[i], &blast[i]);
    parent[bl[i] < 0 ? (n - 1 + ~bl[i]) : bl[i]] = i;
    parent[br[i] < 0 ? (n - 1 + ~br[i]) : br[i]] = i;
  }
  parent[0] = -1;
  // structural check: every node except the root has exactly one parent
  std::vector<int> flags(n, 0);
  for (int i = 0; i < n && !sah; ++i) {
    int p = parent[n - 1 + i];
    while (p >= 0) {
      if (flags[p]++ == 0) break;
      int cs[2] = {bl[p], br[p]};
      for (int k = 0; k < 2; ++k) {
        int c = cs[k];
        const float* lo = c < 0 ? &llo[3 * (size_t)(~c)] : &blo[3 * (size_t)c];
        const float* hi = c < 0 ? &lhi[3 * (size_t)(~c)] : &bhi[3 * (size_t)c];
        for (int a = 0; a < 3; ++a) { blo[3 * p + a] = fminf(blo[3 * p + a], lo[a]); bhi[3 * p + a] = fmaxf(bhi[3 * p + a], hi[a]); }
      }
      p = parent[p];
    }
  }
  std::vector<int> queue(n);
  s->nodes.resize(5 * (size_t)n);
  queue[0] = 0;
  unsigned node_count = 1, tri_count = 0;
  int begin = 0, end = 1;
  s->depth = 0;
  while (begin < end) {
    ++s->depth;
    for (int j = begin; j < end; ++j) {
      int b = queue[j];
      CollapseOut co;
      collapse_select(b, bl.data(), br.data(), bfirst.data(), blast.data(), blo.data(), bhi.data(), llo.data(), lhi.data(), &co);
      uint32_t child_base = node_count; node_count += co.n_inner;
      uint32_t tri_base = tri_count; tri_count += co.n_tris;
      for (int r = 0; r < co.n_inner; ++r) queue[child_base + r] = co.inner_ref[r];
      encode_node8(&blo[3 * (size_t)b], &bhi[3 * (size_t)b], co, child_base, tri_base, &s->nodes[5 * (size_t)j]);
      uint32_t off = 0;
      for (int l = 0; l < co.n_leaf; ++l)
        for (int k = 0; k < co.tri_count[l]; ++k, ++off)
          write_tri(*s, order[co.tri_first[l] + k], &s->tris[12 * (size_t)(tri_base + off)]);
    }
    begin = end; end = (int)node_count;
  }
  s->nodes.resize(5 * (size_t)node_count);
  if (tri_count != (unsigned)n) { delete s; return nullptr; }
  return s;
}

void emu_destroy(EmuScene* s) { delete s; }
int emu_depth(const EmuScene* s) { return s->depth; }
size_t emu_num_nodes(const EmuScene* s) { return s->nodes.size() / 5; }
void emu_get_bvh(const EmuScene* s, void* nodes, void* tris) {
  memcpy(nodes, s->nodes.data(), s->nodes.size() * sizeof(u4));
  memcpy(tris, s->tris.data(), s->tris.size() * sizeof(float));
}

// rays: 8 floats each (ox,oy,oz,tmin,dx,dy,dz,tmax)
void emu_intersect(EmuScene* s, uint32_t domain, const float* rays, size_t n, uint64_t* keys, float* uv,
                   uint64_t* n_nodes, uint64_t* n_tris) {
  uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
  uint64_t tn = 0, tt = 0;
  for (size_t i = 0; i < n; ++i) {
    const float* r = rays + 8 * i;
    RayCtx rc = make_ray_ctx(mk3(r[0], r[1], r[2]), mk3(r[4], r[5], r[6]), r[3]);
    Closest c; c.t = INFINITY; c.u = c.v = 0; c.prim = 0xFFFFFFFFu;
    uint32_t nn = 0, nt = 0;
    bvh8_closest(NodeLoaderH{s->nodes.data()}, TriLoaderH{s->tris.data()}, rc, c, sx, sy, &nn, &nt);
    tn += nn; tt += nt;
    bool hit = c.prim != 0xFFFFFFFFu;
    keys[i] = hit ? make_key(c.t, domain, c.prim) : SPRAY_KEY_MISS;
    if (uv) { uv[2 * i] = hit ? c.u : 0.f; uv[2 * i + 1] = hit ? c.v : 0.f; }
  }
  if (n_nodes) *n_nodes = tn;
  if (n_tris) *n_tris = tt;
}

// MEASUREMENT ONLY (tools/bvh_quality.py): the same BVH8 walked with the hit children of every node visited in
// order of their entry distance (decoded from the quantised planes) instead of the fixed octant order the product
// uses — how many node visits / triangle tests would distance ordering save?
void emu_intersect_sorted(EmuScene* s, const float* rays, size_t n, uint64_t* keys, uint64_t* n_nodes, uint64_t* n_tris) {
  uint64_t tn = 0, tt = 0;
  struct Ent { uint32_t node; float t; };
  std::vector<Ent> stack;
  for (size_t i = 0; i < n; ++i) {
    const float* rr = rays + 8 * i;
    RayCtx r = make_ray_ctx(mk3(rr[0], rr[1], rr[2]), mk3(rr[4], rr[5], rr[6]), rr[3]);
    Closest best; best.t = INFINITY; best.u = best.v = 0; best.prim = 0xFFFFFFFFu;
    stack.clear(); stack.push_back({0u, -INFINITY});
    while (!stack.empty()) {
      Ent e = stack.back(); stack.pop_back();
      if (e.t > best.t) continue;
      const u4* q = &s->nodes[5 * (size_t)e.node];
      ++tn;
#if SPRAY_POP_CULL
      uint32_t unused0, unused1;
      const uint32_t hm = node8_hits(q[0], q[1], q[2], q[3], q[4], r, best.t, &unused0, &unused1);
#else
      const uint32_t hm = node8_hits(q[0], q[1], q[2], q[3], q[4], r, best.t);
#endif
      uint32_t tg = hm & 0x00FFFFFFu;
      const char* lc = getenv("EMU_LEAFCULL");   // measurement: a leaf child's triangles are skipped when its box starts behind the best hit
      const uint32_t metas[2] = {q[1].z, q[1].w};
      while (tg) {
        const int ti = ffs32(tg) - 1; tg &= tg - 1;
        if (lc) {
          int slot = -1;
          for (int sl = 0; sl < 8; ++sl) {
            const uint32_t mb = byte_of(metas[sl >> 2], sl & 3);
            const uint32_t first = mb & 31u, cnt = (mb >> 5) == 1 ? 1u : ((mb >> 5) == 3 ? 2u : ((mb >> 5) == 7 ? 3u : 0u));
            if (mb != 0 && first < 24u && (uint32_t)ti >= first && (uint32_t)ti < first + cnt) slot = sl;
          }
          if (slot >= 0) {
            const uint32_t w = (uint32_t)slot >> 2, j = (uint32_t)slot & 3;
            const uint32_t lo3[3] = {byte_of(w ? q[2].y : q[2].x, j), byte_of(w ? q[2].w : q[2].z, j), byte_of(w ? q[3].y : q[3].x, j)};
            const uint32_t hi3[3] = {byte_of(w ? q[3].w : q[3].z, j), byte_of(w ? q[4].y : q[4].x, j), byte_of(w ? q[4].w : q[4].z, j)};
            const float sc[3] = {u2f((q[0].w & 0xff) << 23), u2f(((q[0].w >> 8) & 0xff) << 23), u2f(((q[0].w >> 16) & 0xff) << 23)};
            const float org[3] = {u2f(q[0].x), u2f(q[0].y), u2f(q[0].z)};
            const float o3[3] = {r.o.x, r.o.y, r.o.z}, id3[3] = {r.idir.x, r.idir.y, r.idir.z};
            float tmin = r.tnear;
            for (int a = 0; a < 3; ++a)
              tmin = fmaxf(tmin, (org[a] + (float)(id3[a] < 0.f ? hi3[a] : lo3[a]) * sc[a] - o3[a]) * id3[a]);
            if (tmin > best.t) continue;
          }
        }
        vec3 v0, e1, e2; uint32_t prim;
        TriLoaderH{s->tris.data()}(q[1].y + (uint32_t)ti, v0, e1, e2, prim);
        ++tt;
        float T, U, V, ad;
        if (tri_test_raw(r.o, r.d, v0, e1, e2, r.tnear, &T, &U, &V, &ad)) {
          const float t = div_rn(T, ad);
          if (t < best.t || (t == best.t && prim < best.prim)) { best.t = t; best.u = div_rn(U, ad); best.v = div_rn(V, ad); best.prim = prim; }
        }
      }
      const uint32_t imask = q[0].w >> 24, ex = q[0].w & 0xff, ey = (q[0].w >> 8) & 0xff, ez = (q[0].w >> 16) & 0xff;
      Ent kids[8]; int nk = 0;
      for (uint32_t bits = hm >> 24; bits; bits &= bits - 1) {
        const uint32_t slot = (uint32_t)(ffs32(bits) - 1) ^ r.octinv;
        const uint32_t rel = (uint32_t)popc32(imask & ~(0xFFFFFFFFu << slot));
        const uint32_t w = slot >> 2, j = slot & 3;
        const uint32_t lo3[3] = {byte_of(w ? q[2].y : q[2].x, j), byte_of(w ? q[2].w : q[2].z, j), byte_of(w ? q[3].y : q[3].x, j)};
        const uint32_t hi3[3] = {byte_of(w ? q[3].w : q[3].z, j), byte_of(w ? q[4].y : q[4].x, j), byte_of(w ? q[4].w : q[4].z, j)};
        const float sc[3] = {u2f(ex << 23), u2f(ey << 23), u2f(ez << 23)};
        const float org[3] = {u2f(q[0].x), u2f(q[0].y), u2f(q[0].z)};
        const float o3[3] = {r.o.x, r.o.y, r.o.z}, id3[3] = {r.idir.x, r.idir.y, r.idir.z};
        float tmin = r.tnear;
        for (int a = 0; a < 3; ++a) {
          const float pl = org[a] + (float)(id3[a] < 0.f ? hi3[a] : lo3[a]) * sc[a];
          tmin = fmaxf(tmin, (pl - o3[a]) * id3[a]);
        }
        kids[nk++] = {q[1].x + rel, tmin};
      }
      const char* ord0 = getenv("EMU_ORDER");
      if (ord0 && !strcmp(ord0, "group")) {      // one distance per node: the nearest hit child's (what a 1-float-per-entry stack could hold)
        float m = INFINITY;
        for (int k = 0; k < nk; ++k) m = fminf(m, kids[k].t);
        for (int k = 0; k < nk; ++k) kids[k].t = m;
      }
      if (ord0 && !strncmp(ord0, "q", 1)) {      // "q<bits>": per-child distance quantised DOWN to <bits> bits of [nearest, farthest hit child]
        const int nb = atoi(ord0 + 1);
        float m = INFINITY;
        for (int k = 0; k < nk; ++k) m = fminf(m, kids[k].t);
        float hi = m;
        for (int k = 0; k < nk; ++k) hi = fmaxf(hi, kids[k].t);
        const float step = (hi - m) / (float)((1 << nb) - 1);
        for (int k = 0; k < nk; ++k) kids[k].t = step > 0.f && step < 1e29f ? m + floorf((kids[k].t - m) / step) * step : m;
      }
      if (ord0 && !strncmp(ord0, "m", 1)) {      // "m<bits>": absolute distance truncated to <bits> mantissa bits (a minifloat byte per child)
        const int nb = atoi(ord0 + 1);
        for (int k = 0; k < nk; ++k) kids[k].t = kids[k].t > 0.f ? u2f(f2u(kids[k].t) & ~((1u << (23 - nb)) - 1u)) : -INFINITY;
      }
      const char* ord = (ord0 && (!strcmp(ord0, "group") || !strncmp(ord0, "q", 1) || !strncmp(ord0, "m", 1))) ? "octant" : ord0;   // "distance": nearest first; "octant": product order, entries culled at pop
      if (!ord || strcmp(ord, "octant"))
        std::sort(kids, kids + nk, [](const Ent& a, const Ent& b) { return a.t > b.t; });   // far first: nearest on top
      for (int k = 0; k < nk; ++k) stack.push_back(kids[k]);
    }
    keys[i] = best.prim != 0xFFFFFFFFu ? make_key(best.t, 0, best.prim) : SPRAY_KEY_MISS;
  }
  *n_nodes = tn; *n_tris = tt;
}

// MEASUREMENT ONLY: a 32-lane model of the product's warp-synchronous loop (traverse.cuh LaneTrav::pass, one
// domain): per pass every lane that wants a node step takes one together, then up to 8 rounds of triangle steps.
// Counts how often each step kind is EXECUTED by the warp and by how many lanes, for two scheduling policies:
//   refill = 0: the product's — a warp takes 32 rays and runs until the longest one ends
//   refill = -K: batches that END once fewer than K lanes are alive; the unfinished rays go back to the queue with
//               their best hit so far and are later restarted from the root (each ray is cut at most once)
//   refill = K: a stream — once K lanes are idle they all take their next rays together (K = 1: the upper bound
//               of what any refill scheme can reach)
// out[0..6] = node executions, node lanes, triangle executions, triangle lanes, passes, refill executions, lanes
void emu_simt(EmuScene* s, const float* rays, size_t n, int refill, uint64_t* out) {
  struct Lane {
    uint32_t ngx, ngy, tgx, tgy; int sp; bool alive; RayCtx r; Closest best;
    size_t ray; bool again;      // ray index; again: this ray was cut once already and must run to its end
    uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
    uint32_t gd0, gd1, s0[SPRAY_STACK_SIZE], s1[SPRAY_STACK_SIZE];   // SPRAY_POP_CULL builds only
  };
  std::vector<Lane> L(32);
  uint64_t ne = 0, nl = 0, te = 0, tl = 0, passes = 0, rex = 0, rl = 0;
  size_t next = 0;
  struct Cut { size_t ray; Closest best; };
  std::vector<Cut> cut;          // refill < 0: rays cut out of a thinning batch, restarted from the root with their best hit
  auto load = [&](Lane& l) {
    if (cut.empty() && next >= n) { l.alive = false; return; }
    l.best.t = INFINITY; l.best.u = l.best.v = 0; l.best.prim = 0xFFFFFFFFu; l.again = false;
    if (!cut.empty()) { l.ray = cut.back().ray; l.best = cut.back().best; l.again = true; cut.pop_back(); }
    else l.ray = next++;
    const float* rr = rays + 8 * l.ray;
    l.r = make_ray_ctx(mk3(rr[0], rr[1], rr[2]), mk3(rr[4], rr[5], rr[6]), rr[3]);
    l.ngx = 0; l.ngy = 0x80000000u; l.tgx = l.tgy = 0; l.sp = 0; l.alive = true; l.gd0 = l.gd1 = 0;
  };
  while (next < n || !cut.empty()) {
    for (auto& l : L) load(l);
    for (;;) {
      int want[32]; bool any = false;
      if (refill < 0) {        // refill = -K: the batch ends once fewer than K lanes are alive; the rest is re-queued
        int alive = 0; bool pinned = false;
        for (auto& l : L) { alive += l.alive; pinned |= l.alive && l.again; }
        if (alive > 0 && alive < -refill && !pinned && next < n) {
          ++rex; rl += alive;
          for (auto& l : L) if (l.alive) { cut.push_back({l.ray, l.best}); l.alive = false; }
          break;
        }
      }
      if (refill > 0) {            // refill = K: all idle lanes take new rays together once at least K lanes are idle
        int idle = 0;
        for (auto& l : L) idle += !l.alive;
        if (idle >= refill && next < n) { ++rex; rl += idle; for (auto& l : L) if (!l.alive) load(l); }
      }
      for (int i = 0; i < 32; ++i) {
        Lane& l = L[i];
        want[i] = 0;
        if (!l.alive) continue;
        if (l.tgy) { want[i] = 3; any = true; continue; }
#if SPRAY_POP_CULL
        while (l.ngy <= 0x00FFFFFFu && l.sp > 0) {
          --l.sp; l.ngx = l.sx[l.sp]; l.gd0 = l.s0[l.sp]; l.gd1 = l.s1[l.sp];
          l.ngy = prune_group(l.sy[l.sp], l.gd0, l.gd1, l.best.t, l.r.octinv);
        }
#else
        if (l.ngy <= 0x00FFFFFFu && l.sp > 0) { --l.sp; l.ngx = l.sx[l.sp]; l.ngy = l.sy[l.sp]; }
#endif
        if (l.ngy > 0x00FFFFFFu) { want[i] = 2; any = true; } else l.alive = false;
      }
      if (!any) { bool left = false; for (auto& l : L) left |= l.alive; if (!left) break; continue; }
      ++passes;
      int c = 0;
      for (int i = 0; i < 32; ++i) if (want[i] == 2) {
        Lane& l = L[i]; ++c;
        const uint32_t imask = l.ngy & 0xffu;
        const int bit = 31 - clz32(l.ngy);
        l.ngy &= ~(1u << bit);
        if (l.ngy > 0x00FFFFFFu && l.sp < SPRAY_STACK_SIZE) { l.sx[l.sp] = l.ngx; l.sy[l.sp] = l.ngy; l.s0[l.sp] = l.gd0; l.s1[l.sp] = l.gd1; ++l.sp; }
        const uint32_t slot = (uint32_t)(bit - 24) ^ l.r.octinv;
        const uint32_t rel = (uint32_t)popc32(imask & ~(0xFFFFFFFFu << slot));
        const u4* q = &s->nodes[5 * (size_t)(l.ngx + rel)];
#if SPRAY_POP_CULL
        const uint32_t hm = node8_hits(q[0], q[1], q[2], q[3], q[4], l.r, l.best.t, &l.gd0, &l.gd1);
#else
        const uint32_t hm = node8_hits(q[0], q[1], q[2], q[3], q[4], l.r, l.best.t);
#endif
        l.ngx = q[1].x; l.tgx = q[1].y;
        l.ngy = (hm & 0xFF000000u) | (q[0].w >> 24);
        l.tgy = hm & 0x00FFFFFFu;
      }
      if (c) { ++ne; nl += c; }
      for (int k = 0; k < 8; ++k) {
        c = 0;
        for (auto& l : L) if (l.alive && l.tgy) {
          ++c;
          const int ti = ffs32(l.tgy) - 1; l.tgy &= l.tgy - 1;
          vec3 v0, e1, e2; uint32_t prim;
          TriLoaderH{s->tris.data()}(l.tgx + (uint32_t)ti, v0, e1, e2, prim);
          float T, U, V, ad;
          if (tri_test_raw(l.r.o, l.r.d, v0, e1, e2, l.r.tnear, &T, &U, &V, &ad)) {
            const float t = div_rn(T, ad);
            if (t < l.best.t || (t == l.best.t && prim < l.best.prim)) {
              l.best.t = t; l.best.prim = prim;
#if SPRAY_POP_CULL
              l.ngy = prune_group(l.ngy, l.gd0, l.gd1, l.best.t, l.r.octinv);
#endif
            }
          }
        }
        if (!c) break;
        ++te; tl += c;
      }
    }
  }
  out[0] = ne; out[1] = nl; out[2] = te; out[3] = tl; out[4] = passes; out[5] = rex; out[6] = rl;
}

void emu_occluded(EmuScene* s, const float* rays, size_t n, uint8_t* occ) {
  uint32_t sx[SPRAY_STACK_SIZE], sy[SPRAY_STACK_SIZE];
  for (size_t i = 0; i < n; ++i) {
    const float* r = rays + 8 * i;
    RayCtx rc = make_ray_ctx(mk3(r[0], r[1], r[2]), mk3(r[4], r[5], r[6]), r[3]);
    occ[i] = bvh8_any(NodeLoaderH{s->nodes.data()}, TriLoaderH{s->tris.data()}, rc, sx, sy, nullptr, nullptr) ? 1 : 0;
  }
}

// closed-form pieces, for cross-checking core.cuh against the oracle
uint32_t emu_rng_init1(uint32_t id) { return rng_init1(id); }
float emu_rng_float(uint32_t* s) { return rng_float(*s); }
uint32_t emu_morton(float x, float y, float z) { return morton30(x, y, z); }
int emu_domain_slab(const float* b, const float* o, const float* d, float* tmin) {
  return domain_slab(b, mk3(o[0], o[1], o[2]), mk3(d[0], d[1], d[2]), SPRAY_RAY_EPS, INFINITY, tmin) ? 1 : 0;
}
// the same test through the ray context's reciprocals (what the fused kernel's top-level walk calls)
int emu_domain_slab_ctx(const float* b, const float* o, const float* d, float* tmin) {
  RayCtx r = make_ray_ctx(mk3(o[0], o[1], o[2]), mk3(d[0], d[1], d[2]), SPRAY_RAY_EPS);
  return domain_slab_inv(b, r.o, exact_inv(r.d.x, r.idir.x), exact_inv(r.d.y, r.idir.y), exact_inv(r.d.z, r.idir.z),
                         SPRAY_RAY_EPS, INFINITY, tmin) ? 1 : 0;
}
uint32_t emu_interp_color(uint32_t c0, uint32_t c1, uint32_t c2, float u, float v) {
  return interp_color(c0, c1, c2, u, v, sub_rn(sub_rn(1.f, u), v));
}
void emu_cosine_hemisphere(float u1, float u2, const float* n, float* wi, float* pdf) {
  vec3 w; cosine_hemisphere(u1, u2, mk3(n[0], n[1], n[2]), &w, pdf); wi[0] = w.x; wi[1] = w.y; wi[2] = w.z;
}
void emu_blinn_phong(float costheta, const float* kd, const float* ks, float sh, const float* li, const float* wi,
                     const float* n, const float* wo, float* out) {
  vec3 r = blinn_phong(costheta, mk3(kd[0], kd[1], kd[2]), mk3(ks[0], ks[1], ks[2]), sh, mk3(li[0], li[1], li[2]),
                       mk3(wi[0], wi[1], wi[2]), mk3(n[0], n[1], n[2]), mk3(wo[0], wo[1], wo[2]));
  out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void emu_camera_dir(const float* cam13, int w, int h, float x, float y, float* dir) {
  CameraDev c;
  for (int k = 0; k < 3; ++k) { c.pos[k] = cam13[k]; c.lowerleft[k] = cam13[3 + k]; c.wvec[k] = cam13[6 + k]; c.hvec[k] = cam13[9 + k]; }
  c.image_w = w; c.image_h = h;
  vec3 d = camera_dir(c, x, y); dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}
}
