"""Known-answer tests that pin the oracle's closed-form pieces against INDEPENDENT restatements written
here in plain Python integers / numpy float32 from the reference source (file:line cited per test).
The reference has no tests of its own: these KATs are what "pins" the oracle."""
import ctypes as C
import json
import math
from collections import OrderedDict

import numpy as np
import pytest

from spray_b200 import scenes

f32 = np.float32
M32 = 0xFFFFFFFF


# ---- RNG: deps/embree/random_sampler.h:34-108 ------------------------------------------------------
def py_murmur_mix(h, k):
    k = (k * 0xcc9e2d51) & M32
    k = ((k << 15) | (k >> 17)) & M32
    k = (k * 0x1b873593) & M32
    h ^= k
    h = ((h << 13) | (h >> 19)) & M32
    return (h * 5 + 0xe6546b64) & M32


def py_murmur_fin(h):
    h ^= h >> 16; h = (h * 0x85ebca6b) & M32
    h ^= h >> 13; h = (h * 0xc2b2ae35) & M32
    h ^= h >> 16
    return h


def py_rng_init1(i):
    return py_murmur_fin(py_murmur_mix(0, i & M32))


def py_rng_float(s):
    s = (s * 1664525 + 1013904223) & M32
    return s, f32(f32(s >> 1) * f32(4.656612873077392578125e-10))


def test_rng_stream(oracle):
    L = oracle.lib()
    # MurmurHash3 fmix of 0 is 0: seed(0) = fin(mix(0,0)); literal pinned from the integer recipe
    assert py_rng_init1(0) == py_murmur_fin(0xe6546b64)
    for seed in [0, 1, 2, 7, 12345, 2 ** 31 - 1, -1, -12345, 1048575 * 3]:
        s_ref = py_rng_init1(seed)
        assert L.orc_rng_init1(seed) == s_ref
        st = C.c_uint32(s_ref)
        for _ in range(16):
            s_ref, f = py_rng_float(s_ref)
            got = L.orc_rng_get_float(C.byref(st))
            assert st.value == s_ref
            assert f32(got) == f and 0.0 <= got < 1.0
    assert L.orc_rng_init2(3, 5) == py_murmur_fin(py_murmur_mix(py_murmur_mix(0, 3), 5))


# ---- Morton: render/morton.h:33-52 -------------------------------------------------------------------
def py_expand(v):
    out = 0
    for b in range(10):
        out |= ((v >> b) & 1) << (3 * b)
    return out


def py_morton(x, y, z):
    q = [int(min(max(f32(c) * f32(1024.0), f32(0.0)), f32(1023.0))) for c in (x, y, z)]
    return py_expand(q[0]) * 4 + py_expand(q[1]) * 2 + py_expand(q[2])


def test_morton(oracle):
    L = oracle.lib()
    assert L.orc_morton(0.0, 0.0, 0.0) == 0
    assert L.orc_morton(1.0, 1.0, 1.0) == 0x3FFFFFFF
    assert L.orc_morton(1.0, 0.0, 0.0) == 0x24924924
    assert L.orc_morton(0.0, 0.0, 1.0) == 0x09249249
    rng = np.random.default_rng(0)
    for x, y, z in rng.uniform(-0.1, 1.1, (200, 3)).astype(np.float32):
        assert L.orc_morton(x, y, z) == py_morton(x, y, z)


# ---- slab test: render/aabb.h:139-169 ----------------------------------------------------------------
def py_slab(b, o, d, t0, t1):
    with np.errstate(all="ignore"):
        inv = [f32(1.0) / f32(x) for x in d]
        s = [int(i < 0) for i in inv]
        lohi = lambda k, a: f32(b[3 * k + a])
        tmin = f32((lohi(s[0], 0) - f32(o[0])) * inv[0]); tmax = f32((lohi(1 - s[0], 0) - f32(o[0])) * inv[0])
        tymin = f32((lohi(s[1], 1) - f32(o[1])) * inv[1]); tymax = f32((lohi(1 - s[1], 1) - f32(o[1])) * inv[1])
        if tmin > tymax or tymin > tmax:
            return False, None
        if tymin > tmin: tmin = tymin
        if tymax < tmax: tmax = tymax
        tzmin = f32((lohi(s[2], 2) - f32(o[2])) * inv[2]); tzmax = f32((lohi(1 - s[2], 2) - f32(o[2])) * inv[2])
        if tmin > tzmax or tzmin > tmax:
            return False, None
        if tzmin > tmin: tmin = tzmin
        if tzmax < tmax: tmax = tzmax
        return bool(tmin < t1 and tmax > t0), tmin


def test_intersect_aabb(oracle):
    L = oracle.lib()
    box = np.array([-10, -10, -10, 10, 9.324713, 10], np.float32)
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    rng = np.random.default_rng(3)
    cases = []
    for _ in range(300):
        o = rng.uniform(-30, 30, 3).astype(np.float32); d = rng.normal(size=3).astype(np.float32)
        cases.append((o, d))
    # origin inside (negative tmin, not clamped), axis-parallel, on-face with zero component (NaN path)
    cases += [(np.array([0, 0, 0], np.float32), np.array([0, 0, 1], np.float32)),
              (np.array([0, 0, 30], np.float32), np.array([0, 0, -1], np.float32)),
              (np.array([10, 0, 30], np.float32), np.array([0, 0, -1], np.float32)),
              (np.array([0, 20, 30], np.float32), np.array([0, 0, -1], np.float32)),
              (np.array([0, 0, -30], np.float32), np.array([0, 0, -1], np.float32))]
    for o, d in cases:
        tmin, tmax = C.c_float(-777), C.c_float(-777)
        got = L.orc_intersect_aabb(fp(box), fp(o), fp(d), f32(1e-3), f32(np.inf), C.byref(tmin), C.byref(tmax))
        exp, etmin = py_slab(box, o, d, f32(1e-3), f32(np.inf))
        assert bool(got) == exp, (o, d)
        if exp:
            assert f32(tmin.value) == etmin
    # inside the box: entry distance is negative
    tmin, tmax = C.c_float(), C.c_float()
    o = np.zeros(3, np.float32); d = np.array([1, 0, 0], np.float32)
    assert L.orc_intersect_aabb(fp(box), fp(o), fp(d), f32(1e-3), f32(np.inf), C.byref(tmin), C.byref(tmax))
    assert tmin.value == -10.0 and tmax.value == 10.0


# ---- colour: util.h:69-98, trimesh_buffer.cc:336-350 ---------------------------------------------------
def test_color_interp_truncates(oracle):
    L = oracle.lib()
    c0, c1, c2 = 0xFF0000, 0x00FF00, 0x0000FF
    assert L.orc_interp_color(c0, c1, c2, 0.0, 0.0) == 0xFF0000
    assert L.orc_interp_color(c0, c1, c2, 1.0, 0.0) == 0x00FF00
    assert L.orc_interp_color(c0, c1, c2, 0.0, 1.0) == 0x0000FF
    # 255 * 0.5 = 127.5 -> truncated to 127 (not rounded)
    assert L.orc_interp_color(c0, c1, c2, 0.5, 0.0) == (127 << 16) | (127 << 8)
    rng = np.random.default_rng(5)
    for _ in range(200):
        cs = [int(x) for x in rng.integers(0, 1 << 24, 3)]
        u = f32(rng.uniform(0, 1)); v = f32(rng.uniform(0, 1 - u))
        w = f32(f32(f32(1.0) - u) - v)
        exp = 0
        for sh in (16, 8, 0):
            ch = [f32((c >> sh) & 0xff) for c in cs]
            s = f32(f32(f32(ch[0] * w) + f32(ch[1] * u)) + f32(ch[2] * v))
            exp |= int(s) << sh
        assert L.orc_interp_color(cs[0], cs[1], cs[2], u, v) == exp
    rgb = np.zeros(3, np.float32)
    L.orc_unpack_color(0x80FF01, rgb.ctypes.data_as(C.POINTER(C.c_float)))
    assert list(rgb) == [f32(128 * 0.00392156862745098), f32(255 * 0.00392156862745098), f32(1 * 0.00392156862745098)]
    assert rgb[1] == f32(1.0)


def test_ppm_quantize(oracle):
    L = oracle.lib()
    px = np.array([[0.0, 0.5, 1.0, 0], [2.0, -1.0, 0.9999, 0]], np.float32)
    out = np.zeros(6, np.uint32)
    L.orc_ppm_quantize(px.ctypes.data_as(C.POINTER(C.c_float)), 2, out.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert list(out) == [0, 511, 1023, 1023, 0, int(f32(0.9999) * f32(1023.0))]


# ---- camera: camera.h:128-209, spray_renderer.inl:87-105 -----------------------------------------------
def test_camera_matches_numpy_restatement(oracle):
    b = np.array([-10, -10, -10, 10, 9.324713, 10], np.float32)
    pos, lookat, up = scenes.default_camera_pose(b)
    p2, l2, u2 = oracle.default_camera_pose(b)
    assert np.array_equal(pos, p2) and np.array_equal(lookat, l2) and np.array_equal(up, u2)
    # default pose: centre + 0.5*|extent| * z
    assert pos[2] == f32(np.sqrt(f32(f32(f32(400.0) + f32(b[4] - b[1]) ** 2) + f32(400.0)), dtype=np.float32) * f32(0.5))
    for (p, l, u, fov, w, h) in [(pos, lookat, up, 90.0, 1024, 1024),
                                 (scenes.WAVELETS64_CAMERA["pos"], scenes.WAVELETS64_CAMERA["lookat"], (0, 1, 0), 90.0, 2048, 2048),
                                 ((3, 4, 5), (0, 1, 0), (0, 1, 0), 45.0, 640, 480)]:
        d = scenes.camera_frame(p, l, u, fov, w, h)
        cam = oracle.make_camera(p, l, u, fov, w, h)
        for k in ("pos", "lowerleft", "wvec", "hvec"):
            assert np.array_equal(np.array(list(getattr(cam, k)), np.float32), d[k]), k
    # fov 90: half height = tan(pi/4) (float) ~ 1 ; the four corner rays of the default camera
    cam = oracle.make_camera(pos, lookat, up, 90.0, 4, 4)
    org = np.zeros(3, np.float32); dr = np.zeros(3, np.float32)
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    oracle.lib().orc_camera_ray(C.byref(cam), f32(0), f32(0), fp(org), fp(dr))
    assert np.array_equal(org, pos)
    s = f32(1.0) / np.sqrt(f32(3.0), dtype=np.float32)
    assert np.allclose(dr, [-s, -s, -s], atol=1e-6)       # pixel (0,0) = lower-left CORNER (no +0.5)
    oracle.lib().orc_camera_ray(C.byref(cam), f32(2), f32(2), fp(org), fp(dr))
    assert np.allclose(dr, [0, 0, -1], atol=1e-7)         # image centre looks down -z


# ---- tiles: tile.cc:204-235, 317-391 ---------------------------------------------------------------------
@pytest.mark.parametrize("w,h,spp,nr,cap", [(1024, 1024, 1, 1, 1 << 20), (2048, 2048, 1, 1, 1 << 20),
                                            (2048, 2048, 1, 2, 1 << 20), (512, 512, 1, 2, 1 << 20),
                                            (4096, 4096, 4, 8, 1 << 20), (400, 400, 4, 3, 1 << 20),
                                            (100, 37, 2, 3, 1000), (2048, 2048, 1, 1, 1 << 30)])
def test_image_schedule_tiles(oracle, w, h, spp, nr, cap):
    covered = np.zeros((h, w), np.int32)
    for rank in range(nr):
        got = oracle.image_schedule_tiles(w, h, spp, nr, rank, cap)
        assert got == scenes.image_schedule_tiles(w, h, spp, nr, rank, cap)
        for (x, y, tw, th) in got:
            covered[y:y + th, x:x + tw] += 1
    assert (covered == 1).all()       # stripes x tiles tile the image exactly once
    if (w, h, spp, nr) == (2048, 2048, 1, 1):
        assert oracle.image_schedule_tiles(w, h, spp, nr, 0, 1 << 20) == [(0, 512 * i, 2048, 512) for i in range(4)]


# ---- LRU: lru_cache.cc:65-171 ------------------------------------------------------------------------------
def test_lru_sequence(oracle):
    rng = np.random.default_rng(11)
    for nd, cap in [(8, 3), (64, 16), (5, 5), (4, 1)]:
        req = rng.integers(0, nd, 400).astype(np.int32)
        hits = np.zeros(len(req), np.int32); blocks = np.zeros(len(req), np.int32)
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))
        oracle.lib().orc_lru_sequence(nd, cap, ip(req), len(req), ip(hits), ip(blocks))
        od = OrderedDict()
        for k, d in enumerate(req):
            d = int(d)
            if d in od:
                od.move_to_end(d); assert hits[k] == 1 and blocks[k] == od[d]
            else:
                assert hits[k] == 0
                if len(od) < cap:
                    od[d] = len(od)
                else:
                    _, blk = od.popitem(last=False); od[d] = blk
                assert blocks[k] == od[d]


# ---- partition: data_partition.h:59-137 ----------------------------------------------------------------------
def test_partition_wavelets64(oracle):
    sc = scenes.wavelets64()
    wb = np.stack([d["bounds"] for d in sc["domains"]]).astype(np.float32)
    sb = scenes.scene_bounds(sc)
    for nranks in (1, 2, 4, 8):
        d2r = np.zeros(64, np.int32); codes = np.zeros(64, np.uint32)
        oracle.lib().orc_partition(64, wb.ctypes.data_as(C.POINTER(C.c_float)), sb.ctypes.data_as(C.POINTER(C.c_float)),
                                   nranks, d2r.ctypes.data_as(C.POINTER(C.c_int32)), codes.ctypes.data_as(C.POINTER(C.c_uint32)))
        assert np.bincount(d2r, minlength=nranks).tolist() == [64 // nranks] * nranks
        assert len(set(codes.tolist())) == 64
        # contiguous Morton runs: rank is monotone in code order
        order = np.argsort(codes, kind="stable")
        assert (np.diff(d2r[order]) >= 0).all()
        import partition_ref as partition
        assert partition.insitu_partition(wb, sb, nranks).tolist() == d2r.tolist()


# ---- scene description vs the reference's own wavelets64.spray ---------------------------------------------------
def test_wavelets64_scene_matches_reference_file():
    ref = scenes.scene_from_golden_json()
    sc = scenes.wavelets64()
    assert len(ref["domains"]) == 64 and len(ref["lights"]) == 1
    assert ref["lights"][0] == {"type": "point", "v": [0.0, 500.0, 1000.0, 1.0, 1.0, 1.0]}
    for i, (r, d) in enumerate(zip(ref["domains"], sc["domains"])):
        assert r["file"] == "wavelet.ply" and r["face"] == 5480 and r["vertex"] == 2840
        assert np.allclose(r["translate"], d["translate"], atol=5e-7), i   # the file prints 6 decimals
        ob = np.array(r["bound"], np.float32)
        m = None if not any(r["translate"]) else scenes.translate_mat4(np.array(r["translate"], np.float32))
        assert np.array_equal(scenes.world_bounds(ob, m), d["bounds"]), i
    v, f, c = scenes.load_wavelet_mesh()
    assert v.shape == (2840, 3) and f.shape == (5480, 3) and c.shape == (2840,)
    assert np.array_equal(np.concatenate([v.min(0), v.max(0)]), np.array([-10, -10, -10, 10, 9.324713, 10], np.float32))
    assert len(np.unique(c)) == 221 and f.max() == 2839


# ---- sampling / shading closed forms: sampler.h:49-128, reflection.h:202-214 ------------------------------------------
def test_cosine_hemisphere_and_blinn(oracle):
    L = oracle.lib()
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    rng = np.random.default_rng(2)
    for _ in range(100):
        n = rng.normal(size=3); n = (n / np.linalg.norm(n)).astype(np.float32)
        u1, u2 = f32(rng.random()), f32(rng.random())
        wi = np.zeros(3, np.float32); pdf = C.c_float()
        L.orc_cosine_hemisphere(u1, u2, fp(n), fp(wi), C.byref(pdf))
        assert abs(np.linalg.norm(wi) - 1) < 1e-5
        cosn = float(np.dot(wi, n))
        assert cosn > -1e-4 and abs(pdf.value - max(cosn, 0) / math.pi) < 2e-4
    # centre of the square maps to the pole: wi == n, pdf = 1/pi
    n = np.array([0, 0, 1], np.float32); wi = np.zeros(3, np.float32); pdf = C.c_float()
    L.orc_cosine_hemisphere(f32(0.5), f32(0.5), fp(n), fp(wi), C.byref(pdf))
    assert np.allclose(wi, n, atol=1e-7) and abs(pdf.value - 1 / math.pi) < 1e-7
    # blinnPhong with wi = wo = n: li * (kd*cos + ks * 1^shin)
    out = np.zeros(3, np.float32)
    kd = np.array([0.2, 0.4, 0.6], np.float32); ks = np.array([0.4, 0.4, 0.4], np.float32); li = np.ones(3, np.float32)
    L.orc_blinn_phong(f32(1.0), fp(kd), fp(ks), f32(10.0), fp(li), fp(n), fp(n), fp(n), fp(out))
    assert np.allclose(out, kd + ks, atol=1e-7)


def test_tri_test_conventions(oracle):
    """u,v are relative to v1,v2 (w = 1-u-v weights v0): rays.h:246-274 + trimesh_buffer.cc:343-350."""
    L = oracle.lib()
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    v0 = np.array([0, 0, 0], np.float32); v1 = np.array([1, 0, 0], np.float32); v2 = np.array([0, 1, 0], np.float32)
    d = np.array([0, 0, -1], np.float32)
    hit, t, u, v = C.c_int32(), C.c_float(), C.c_float(), C.c_float()
    for (px, py) in [(0.25, 0.5), (0.1, 0.1), (0.7, 0.2)]:
        o = np.array([px, py, 2], np.float32)
        L.orc_tri_test(fp(o), fp(d), fp(v0), fp(v1), fp(v2), f32(1e-3), C.byref(hit), C.byref(t), C.byref(u), C.byref(v))
        assert hit.value == 1 and abs(t.value - 2) < 1e-6 and abs(u.value - px) < 1e-6 and abs(v.value - py) < 1e-6
    # no back-face culling; inclusive edges; t must exceed tnear
    o = np.array([0.25, 0.25, -2], np.float32); d2 = np.array([0, 0, 1], np.float32)
    L.orc_tri_test(fp(o), fp(d2), fp(v0), fp(v1), fp(v2), f32(1e-3), C.byref(hit), C.byref(t), C.byref(u), C.byref(v))
    assert hit.value == 1
    o = np.array([0.5, 0.0, 2], np.float32)
    L.orc_tri_test(fp(o), fp(d), fp(v0), fp(v1), fp(v2), f32(1e-3), C.byref(hit), C.byref(t), C.byref(u), C.byref(v))
    assert hit.value == 1 and v.value == 0.0
    o = np.array([0.25, 0.25, 0.0005], np.float32)
    L.orc_tri_test(fp(o), fp(d), fp(v0), fp(v1), fp(v2), f32(1e-3), C.byref(hit), C.byref(t), C.byref(u), C.byref(v))
    assert hit.value == 0
    o = np.array([0.6, 0.6, 2], np.float32)
    L.orc_tri_test(fp(o), fp(d), fp(v0), fp(v1), fp(v2), f32(1e-3), C.byref(hit), C.byref(t), C.byref(u), C.byref(v))
    assert hit.value == 0


def test_philox4x32_10_known_answers(oracle):
    """Philox-4x32-10 (Salmon et al., SC'11) as the soup generator of bench configs C3 / C4 uses it: the three known-answer
    vectors published with Random123 (kat_vectors: counter / key all zero, all ones, digits of pi)."""
    kat = [([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
           ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
           ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0], [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1])]
    for ctr, key, want in kat:
        assert [int(x) for x in oracle.philox4x32_10(ctr, key)] == want


def test_soup_generator_properties(oracle):
    """orc_gen_soup: deterministic, a function of (key, primID) only (a prefix of a larger soup with the same h is the
    smaller soup), inside its cell, edge lengths in [0.5, 2] h before clipping, directions spread over the sphere."""
    lo = np.array([20, 0, 40], np.float32); hi = lo + np.float32(20)
    n = 1 << 14
    h = oracle.soup_h(n, lo, hi)
    v, f, c = oracle.gen_soup(0x5EED0003, n, lo, hi, h)
    v2, _, _ = oracle.gen_soup(0x5EED0003, n, lo, hi, h)
    assert np.array_equal(v, v2) and np.array_equal(f.reshape(-1), np.arange(3 * n, dtype=np.uint32))
    vp, _, _ = oracle.gen_soup(0x5EED0003, n // 4, lo, hi, h)
    assert np.array_equal(vp, v[:3 * (n // 4)])
    vo, _, _ = oracle.gen_soup(0x5EED0004, n, lo, hi, h)
    assert not np.array_equal(vo, v)
    assert (v >= lo).all() and (v <= hi).all() and v.min(0).max() - lo.max() < 1 and (c >> 24 == 0).all()
    t = v.reshape(n, 3, 3)
    inner = ((t > lo + 2.5 * h) & (t < hi - 2.5 * h)).all(axis=(1, 2))              # triangles no clipping touched
    e1 = t[inner, 1] - t[inner, 0]
    ln = np.linalg.norm(e1, axis=1)
    assert inner.sum() > n // 3 and ln.min() >= 0.4999 * h and ln.max() <= 2.0001 * h
    d = e1 / ln[:, None]
    assert np.abs(d.mean(0)).max() < 0.03 and np.abs((d * d).mean(0) - 1 / 3).max() < 0.02
