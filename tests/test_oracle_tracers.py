"""Oracle self-consistency: the FP32 BVH tracer (O2) must equal FP32 brute force (O2b) bit for bit, the
double-precision truth (O1) must equal its own brute force, and O2 must agree with O1 on every ray O1
does not flag as near-tie / edge-graze.  Also the rendered image sanity for config C1/C2 shapes."""
import numpy as np
import pytest

from conftest import random_rays
from spray_b200 import scenes


@pytest.fixture(scope="module")
def c1(oracle):
    sc = scenes.wavelet_single()
    return sc, oracle.Scene(sc["domains"])


@pytest.fixture(scope="module")
def c2(oracle):
    sc = scenes.wavelets64()
    return sc, oracle.Scene(sc["domains"])


def _cam(oracle, sc, w, h, pose=None):
    if pose is None:
        pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    else:
        pos, lookat, up = pose
    return oracle.camera_from_desc(scenes.camera_frame(pos, lookat, up, 90.0, w, h))


def test_c1_bvh_equals_brute_and_truth(oracle, c1):
    sc, osc = c1
    cam = _cam(oracle, sc, 128, 128)
    rays = np.concatenate([oracle.gen_eye_rays(cam, (0, 0, 128, 128)), random_rays(6000, [-10] * 3, [10] * 3, seed=1)])
    k, uv = osc.intersect(rays)
    kb, uvb = osc.intersect(rays, brute=True)
    assert np.array_equal(k, kb) and np.array_equal(uv, uvb)
    dom, prim, t, fl = osc.intersect_f64(rays, use_bvh=True)
    dom2, prim2, t2, fl2 = osc.intersect_f64(rays, use_bvh=False)
    assert np.array_equal(dom, dom2) and np.array_equal(prim, prim2) and np.array_equal(t, t2) and np.array_equal(fl, fl2)
    tf, df, pf = oracle.key_fields(k)
    hit = k != oracle.KEY_MISS
    clean = fl == 0
    both = hit & (dom >= 0)
    assert (hit == (dom >= 0))[clean].mean() > 0.999
    assert np.array_equal(pf[both & clean], prim[both & clean])
    assert (np.abs(tf[both] - t[both]) <= 1e-5 * t[both]).all()
    assert (pf[both] == prim[both]).mean() >= 0.95


def test_c2_multidomain_consistency(oracle, c2):
    sc, osc = c2
    p = scenes.WAVELETS64_CAMERA
    cam = _cam(oracle, sc, 96, 96, (p["pos"], p["lookat"], p["up"]))
    rays = oracle.gen_eye_rays(cam, (0, 0, 96, 96))
    k, uv = osc.intersect(rays)
    kb, uvb = osc.intersect(rays, brute=True)
    assert np.array_equal(k, kb) and np.array_equal(uv, uvb)
    counts, ids, tm = osc.domain_lists(rays, cap=32)
    assert counts.max() <= 10 and counts.min() >= 0   # 4x4x4 grid: at most 4+4+4-2 cells on a line
    for i in np.random.default_rng(0).integers(0, len(rays), 200):
        c = counts[i]
        pairs = list(zip(tm[i, :c].tolist(), ids[i, :c].tolist()))
        assert pairs == sorted(pairs)                 # (tmin, id) order, rays.h:71-79
    # per-domain queries min-merged over the entered domains == the all-domain query
    best = np.full(len(rays), oracle.KEY_MISS, np.uint64)
    for d in range(64):
        sel = (ids == d).any(1)
        if sel.any():
            kd, _ = osc.intersect_domain(d, rays[sel])
            best[sel] = np.minimum(best[sel], kd)
    assert np.array_equal(best, k)
    dom, prim, t, fl = osc.intersect_f64(rays)
    tf, df, pf = oracle.key_fields(k)
    both = (k != oracle.KEY_MISS) & (dom >= 0)
    clean = both & (fl == 0)
    assert np.array_equal(df[clean], dom[clean]) and np.array_equal(pf[clean], prim[clean])
    assert ((df[both] == dom[both]) & (pf[both] == prim[both])).mean() >= 0.95


def test_occluded_matches_intersect(oracle, c2):
    sc, osc = c2
    b = scenes.scene_bounds(sc)
    rays = random_rays(4000, b[:3], b[3:], seed=4, spread=0.3)
    occ = osc.occluded(rays)
    k, _ = osc.intersect(rays)
    assert np.array_equal(occ.astype(bool), k != oracle.KEY_MISS)


def test_normals_follow_reference_loop(oracle, c1):
    """computeNormals (trimesh_buffer.cc:267-326) restated in numpy with np.add.at in face order."""
    sc, osc = c1
    d = sc["domains"][0]
    v, f = d["verts"], d["faces"].astype(np.int64)
    n = np.zeros_like(v)
    u = (v[f[:, 1]] - v[f[:, 0]]).astype(np.float32); w = (v[f[:, 2]] - v[f[:, 0]]).astype(np.float32)
    fn = np.stack([(u[:, 1] * w[:, 2]).astype(np.float32) - (u[:, 2] * w[:, 1]).astype(np.float32),
                   (u[:, 2] * w[:, 0]).astype(np.float32) - (u[:, 0] * w[:, 2]).astype(np.float32),
                   (u[:, 0] * w[:, 1]).astype(np.float32) - (u[:, 1] * w[:, 0]).astype(np.float32)], 1).astype(np.float32)
    for i in range(len(f)):      # serial, face order = summation order
        for k in range(3):
            n[f[i, k]] = (n[f[i, k]] + fn[i]).astype(np.float32)
    assert np.array_equal(osc.normals(0), n)


def test_render_c1_small(oracle, c1):
    sc, osc = c1
    W = H = 96
    cam = _cam(oracle, sc, W, H)
    sh, keep = oracle.make_shade(bounces=1, pixel_samples=1, lights=sc["lights"], **scenes.WAVELETS64_BLINN)
    tiles = scenes.image_schedule_tiles(W, H, 1, 1, 0)
    fb, st = osc.render(cam, sh, tiles)
    assert st.eye_rays == W * H and st.radiance_rays == W * H
    rays = oracle.gen_eye_rays(cam, (0, 0, W, H))
    k, _ = osc.intersect(rays)
    nhit = int((k != oracle.KEY_MISS).sum())
    assert 0 < st.shadow_rays <= nhit
    img = fb[..., :3]
    lit = (img.sum(-1) > 0)
    assert lit.sum() > 0.05 * W * H and img.max() <= 1.5 and img.min() >= 0
    assert not lit.reshape(-1)[k == oracle.KEY_MISS].any()       # background stays black
    # bounces=2 adds indirect light only where there was a hit; image changes but stays finite
    sh2, keep2 = oracle.make_shade(bounces=2, pixel_samples=1, lights=sc["lights"], **scenes.WAVELETS64_BLINN)
    fb2, st2 = osc.render(cam, sh2, tiles)
    assert st2.radiance_rays > st.radiance_rays and np.isfinite(fb2).all()
    assert (fb2[..., :3] >= fb[..., :3] - 1e-6).all()
