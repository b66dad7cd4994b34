"""GPU parity at the sizes bench.py measures (pytest -m gpu; VERDICT r1 "What's weak" 1b).

Round 1 compared nothing larger than a 5 480-triangle domain with the oracle, yet benched 64 x 1 Mi-triangle
domains (C3), the 8 Mi-triangle commit (C4) and their out-of-core sweep.  Here:

* `spraygpu_scene_commit` of ONE 1 Mi- and ONE 8 Mi-triangle soup domain (the two-level digit scan and the 256+
  sort tiles of build.cu) followed by random + eye rays: keys / (u,v) / occlusion bit-exact against the oracle's
  FP32 BVH tracer O2 (own binned-SAH tree, same S1 arithmetic) and, on a subsample, against the FP32 brute force O2b;
* a C3-shaped render (2x2x2 domains x 128 Ki triangles, bounces 3, 1 point light): ray counts equal, RMSE <= 1e-4;
* the same scene through the out-of-core sweep with 2 and 5 cache blocks;
* adversarial trees for the traversal stacks: 10^5 triangles with ONE centroid (identical Morton keys: the
  LBVH is built from the index tie-break alone), a long chain of nested triangles, and 512 nested domain boxes.

All through the C ABI (spray_b200.capi -> libspraygpu.so)."""
import numpy as np
import pytest

from conftest import make_rays, random_rays
from test_hostemu import adversarial_meshes
from spray_b200 import capi, scenes

pytestmark = pytest.mark.gpu
RMSE_TOL = 1e-4


def _rmse(a, b):
    return float(np.sqrt(np.mean((np.asarray(a, np.float64)[..., :3] - np.asarray(b, np.float64)[..., :3]) ** 2)))


def _one_domain(oracle, dev, ntris, seed, n_random, eye, n_brute):
    lo, hi = (0.0, 0.0, 0.0), (20.0, 20.0, 20.0)
    v, f, rgb = scenes.soup_domain(seed, ntris, lo, hi)
    dom = dict(verts=v, faces=f, rgb=rgb, transform=None, bounds=np.array(lo + hi, np.float32))
    osc = oracle.Scene([dom])
    s = dev.new_scene()
    s.set_mesh(v, f, rgb)
    s.commit()
    inf = s.info()
    assert inf.num_faces == ntris and 0 < inf.num_nodes < ntris and 0 < inf.bvh_depth <= 48
    assert np.array_equal(np.array(list(inf.bounds), np.float32), np.concatenate([v.min(0), v.max(0)]))
    sc = dict(domains=[dom], lights=[])
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cam = oracle.camera_from_desc(scenes.camera_frame(pos, lookat, up, 90.0, eye, eye))
    rays = np.concatenate([oracle.gen_eye_rays(cam, (0, 0, eye, eye)), random_rays(n_random, lo, hi, seed=seed + 1),
                           random_rays(n_random // 4, lo, hi, seed=seed + 2, spread=-0.3)])     # origins inside
    n = len(rays)
    keys = np.zeros(n, np.uint64); uv = np.zeros((n, 2), np.float32); occ = np.zeros(n, np.uint8)
    s.intersect(0, rays, n, keys, uv)
    s.occluded(rays, n, occ)
    ko, uvo = osc.intersect_domain(0, rays)
    assert (ko != oracle.KEY_MISS).mean() > 0.5
    assert np.array_equal(keys, ko)
    assert np.array_equal(uv, uvo)
    assert np.array_equal(occ, osc.occluded_domain(0, rays))
    sub = np.random.default_rng(seed).choice(n, n_brute, replace=False)
    kb, uvb = osc.intersect(rays[sub], brute=True)                   # O2b: every triangle, no tree at all
    assert np.array_equal(keys[sub], kb) and np.array_equal(uv[sub], uvb)
    # normals (a5) at this size: the (vertex, face) pair sort runs 3 * ntris keys through the same radix sort
    assert np.array_equal(s.normals(), osc.normals(0))
    s.close()


def test_commit_1Mi_triangles_vs_oracle(oracle):
    dev = capi.Device(0)
    _one_domain(oracle, dev, 1 << 20, seed=11, n_random=100000, eye=256, n_brute=1500)
    dev.close()


def test_commit_8Mi_triangles_vs_oracle(oracle):
    """SPRAYGPU_MAX_PRIMS-sized domain (C4's per-domain size, the 23-bit primID field filled)."""
    dev = capi.Device(0)
    _one_domain(oracle, dev, 1 << 23, seed=12, n_random=40000, eye=128, n_brute=256)
    dev.close()


@pytest.fixture(scope="module")
def c3s():
    """C3's shape at a size the oracle renders in seconds: 2x2x2 soup domains x 128 Ki triangles."""
    return scenes.soup_scene((2, 2, 2), 1 << 17)


def _c3_camera(W):
    return scenes.camera_frame((55.0, 50.0, 47.0), (20.0, 20.0, 20.0), (0, 1, 0), 90.0, W, W)


def test_render_c3_shape(oracle, c3s):
    """bounces 3 (primary + 2 indirect generations + shadow rays), incoherent rays in 1 Mi triangles."""
    sc = c3s
    osc = oracle.Scene(sc["domains"])
    W = 224
    cd = _c3_camera(W)
    kw = dict(bounces=3, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    tiles = scenes.image_schedule_tiles(W, W, 1, 1, 0, 1 << 14)
    sh_o, keep = oracle.make_shade(lights=sc["lights"], **kw)
    fb_o, st_o = osc.render(oracle.camera_from_desc(cd), sh_o, tiles)
    dev = capi.Device(0)
    dev.load_scene(sc)
    cam, sh = capi.make_camera(cd), capi.make_shade(lights=sc["lights"], **kw)
    fb = np.zeros_like(fb_o)
    tot = {}
    for t in tiles:
        for k, v in dev.trace_tile(cam, sh, t, fb, stats=True).as_dict().items():
            tot[k] = tot.get(k, 0) + v
    assert tot["eye_rays"] == st_o.eye_rays == W * W
    assert tot["radiance_rays"] == st_o.radiance_rays > 1.3 * W * W    # most hits spawn bounce rays
    assert tot["shadow_rays"] == st_o.shadow_rays
    assert _rmse(fb, fb_o) <= RMSE_TOL
    # the frame call (one persistent launch) and closest-hit keys of the eye rays
    import torch
    fb2 = torch.zeros((W, W, 4), dtype=torch.float32, device="cuda")
    st2 = dev.trace_frame(cam, sh, tiles, fb2, stats=True).as_dict()
    assert st2["radiance_rays"] == st_o.radiance_rays and st2["shadow_rays"] == st_o.shadow_rays
    assert _rmse(fb2.cpu().numpy(), fb_o) <= RMSE_TOL
    rays = oracle.gen_eye_rays(oracle.camera_from_desc(cd), (0, 0, W, W))
    keys = np.zeros(len(rays), np.uint64); uv = np.zeros((len(rays), 2), np.float32)
    dev.intersect_domains(rays, len(rays), keys, uv)
    ko, uvo = osc.intersect(rays)
    assert np.array_equal(keys, ko) and np.array_equal(uv, uvo)
    dev.close()


@pytest.mark.parametrize("capacity", [2, 5])
def test_ooc_sweep_c3_shape(oracle, c3s, capacity):
    """C4's mechanism (bench.py --workload c4s) at this size: device LRU smaller than the scene."""
    import torch
    sc = c3s
    W = 160
    cd = _c3_camera(W)
    kw = dict(bounces=2, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    tiles = [(0, 0, W, W)]
    osc = oracle.Scene(sc["domains"])
    sh_o, keep = oracle.make_shade(lights=sc["lights"], **kw)
    fb_o, st_o = osc.render(oracle.camera_from_desc(cd), sh_o, tiles)
    dev = capi.Device(0)
    cache = capi.LruDomainCache(dev, sc, capacity)
    fb = torch.zeros((W, W, 4), dtype=torch.float32, device="cuda")
    st = dev.trace_tile_ooc(capi.make_camera(cd), capi.make_shade(lights=sc["lights"], **kw), tiles[0], fb, capacity,
                            cache.load, stats=True).as_dict()
    dev.synchronize()
    assert st["radiance_rays"] == st_o.radiance_rays and st["shadow_rays"] == st_o.shadow_rays
    assert _rmse(fb.cpu().numpy(), fb_o) <= RMSE_TOL
    assert cache.misses > capacity                                      # domains were evicted and re-built
    dev.close()


def test_adversarial_trees_do_not_overflow_the_stacks(oracle):
    """The traversal stacks (48 BVH8 levels, 24 top-level entries) trap on overflow instead of dropping a subtree;
    commit refuses trees deeper than the stack.  These inputs are the deepest trees the builder can be made to
    produce; every result must still equal the oracle."""
    dev = capi.Device(0)
    # (1) 10^5 triangles sharing ONE box centre (identical Morton keys), (2) a chain of nested shrinking triangles
    for verts, faces in adversarial_meshes():
        lo, hi = verts.min(0) - 1, verts.max(0) + 1
        dom = dict(verts=verts, faces=faces, rgb=None, transform=None, bounds=np.concatenate([lo, hi]).astype(np.float32))
        osc = oracle.Scene([dom])
        sc = dev.new_scene(); sc.set_mesh(verts, faces, None); sc.commit()
        assert 0 < sc.info().bvh_depth <= 48
        rays = random_rays(30000, lo, hi, seed=3)
        keys = np.zeros(len(rays), np.uint64); occ = np.zeros(len(rays), np.uint8)
        sc.intersect(0, rays, len(rays), keys); sc.occluded(rays, len(rays), occ)
        assert np.array_equal(keys, osc.intersect_domain(0, rays)[0])
        assert np.array_equal(occ, osc.occluded_domain(0, rays))
        sc.close()
    # (3) 512 nested domain boxes (every ray through the centre enters all of them): top-level stack + domain lists
    nd = 512
    tv, tf, trgb = scenes.soup_domain(1, 64, (-1, -1, -1), (1, 1, 1))
    doms = []
    for i in range(nd):
        h = np.float32(1.0 + 0.05 * i)
        doms.append(dict(verts=(tv * h).astype(np.float32), faces=tf, rgb=trgb, transform=None,
                         bounds=np.array([-h, -h, -h, h, h, h], np.float32)))
    scn = dict(domains=doms, lights=[(scenes.LIGHT_POINT, (0.0, 500.0, 1000.0), (1.0, 1.0, 1.0))])
    osc = oracle.Scene(doms)
    dev.load_scene(scn)
    rays = np.concatenate([random_rays(20000, [-30] * 3, [30] * 3, seed=8, spread=0.2),
                           random_rays(5000, [-2] * 3, [2] * 3, seed=9, spread=-0.5)])
    keys = np.zeros(len(rays), np.uint64); uv = np.zeros((len(rays), 2), np.float32); occ = np.zeros(len(rays), np.uint8)
    dev.intersect_domains(rays, len(rays), keys, uv)
    dev.occluded_domains(rays, len(rays), occ)
    ko, uvo = osc.intersect(rays)
    assert np.array_equal(keys, ko) and np.array_equal(uv, uvo)
    assert np.array_equal(occ, osc.occluded(rays))
    cap = 520
    sub = rays[:2000]
    counts = np.zeros(len(sub), np.int32); ids = np.full((len(sub), cap), -1, np.int32); tm = np.zeros((len(sub), cap), np.float32)
    dev.domain_lists(sub, len(sub), cap, counts, ids, tm)
    co, io, to = osc.domain_lists(sub, cap)
    assert np.array_equal(counts, co) and counts.max() == nd
    mask = np.arange(cap)[None, :] < counts[:, None]
    assert np.array_equal(ids[mask], io[mask]) and np.array_equal(tm[mask], to[mask])
    # and a render through the fused kernel (top-level BVH walked lazily from shared memory)
    W = 96
    cd = scenes.camera_frame((40.0, 35.0, 50.0), (0.0, 0.0, 0.0), (0, 1, 0), 60.0, W, W)
    kw = dict(bounces=2, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    sh_o, keep = oracle.make_shade(lights=scn["lights"], **kw)
    fb_o, st_o = osc.render(oracle.camera_from_desc(cd), sh_o, [(0, 0, W, W)])
    fb = np.zeros_like(fb_o)
    st = dev.trace_tile(capi.make_camera(cd), capi.make_shade(lights=scn["lights"], **kw), (0, 0, W, W), fb, stats=True).as_dict()
    assert st["radiance_rays"] == st_o.radiance_rays and st["shadow_rays"] == st_o.shadow_rays
    assert _rmse(fb, fb_o) <= RMSE_TOL
    dev.close()


def test_glass_deep_bounces_queue_bound(oracle):
    """both secondary rays of a glass hit are pushed; the per-warp queue is sized from `bounces`
    (63 + 32 * (bounces - 1) + 32 entries) instead of trapping at bounces >= 4.  bounces 7 here."""
    sc = scenes.wavelets64(8)
    for i, d in enumerate(sc["domains"]):
        d["mtl"] = (scenes.MTL_GLASS, (1.0, 1.5, 0)) if i % 2 == 0 else (scenes.MTL_DIFFUSE, (1, 1, 1))
    osc = oracle.Scene(sc["domains"])
    W = 96
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cd = scenes.camera_frame(pos + np.array([30, 20, 10], np.float32), lookat, up, 70.0, W, W)
    kw = dict(bounces=7, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    sh_o, keep = oracle.make_shade(lights=sc["lights"], **kw)
    fb_o, st_o = osc.render(oracle.camera_from_desc(cd), sh_o, [(0, 0, W, W)])
    dev = capi.Device(0); dev.load_scene(sc)
    fb = np.zeros_like(fb_o)
    st = dev.trace_tile(capi.make_camera(cd), capi.make_shade(lights=sc["lights"], **kw), (0, 0, W, W), fb, stats=True).as_dict()
    # the oracle's per-sample chain stack holds 8 rays (spray_oracle.c trace_sample): equal counts mean neither side dropped one
    assert st["radiance_rays"] == st_o.radiance_rays > 2 * W * W // 3
    assert st["shadow_rays"] == st_o.shadow_rays
    assert _rmse(fb, fb_o) <= RMSE_TOL
    dev.close()


def test_set_mesh_rejects_out_of_range_indices():
    """a malformed PLY must not become out-of-bounds device writes at commit."""
    dev = capi.Device(0)
    v, f, rgb = scenes.soup_domain(2, 1000, (0, 0, 0), (1, 1, 1))
    bad = f.copy(); bad[777, 1] = len(v)            # one past the end
    s = dev.new_scene()
    with pytest.raises(capi.SprayGpuError, match="out of range"):
        s.set_mesh(v, bad, rgb)
    with pytest.raises(capi.SprayGpuError):
        s.commit()
    s.set_mesh(v, f, rgb); s.commit()               # the scene object stays usable
    assert s.info().num_faces == 1000
    # NaN vertices: commit must not fault (collapse_select's slot assignment guard)
    vn = v.copy(); vn[5] = np.nan
    s.set_mesh(vn, f, rgb); s.commit()
    rays = random_rays(2000, [0] * 3, [1] * 3, seed=1)
    keys = np.zeros(len(rays), np.uint64)
    s.intersect(0, rays, len(rays), keys)
    s.close(); dev.close()


def test_soup_generator_device_equals_oracle_bit_for_bit(oracle):
    """spraygpu_gen_soup (the C3 / C4 bench scene, generated on the device) against the CPU restatement: identical
    vertices, indices and colours for several keys / cells / sizes up to the benched 2^20 triangles per domain — the
    benched scene IS reproducible on the oracle — and a scene built from it traces to the oracle's keys."""
    import torch
    for key, n, lo in ((0x5EED0000, 1, (0, 0, 0)), (0x5EED0007, 1000, (20, 40, 0)), (0x5EED003F, 1 << 16, (60, 60, 60)),
                       (0x5EED0015, 1 << 20, (20, 0, 40))):
        lo = np.array(lo, np.float32); hi = lo + np.float32(20)
        v, f, c = scenes.soup_domain_device(key - 0x5EED0000, n, lo.tolist(), hi.tolist())
        torch.cuda.synchronize()
        vo, fo, co = oracle.gen_soup(key, n, lo, hi, scenes.soup_h(n, lo, hi))
        assert np.array_equal(v.cpu().numpy().view(np.uint32), vo.view(np.uint32)), (hex(key), n)
        assert np.array_equal(f.cpu().numpy().view(np.uint32), fo) and np.array_equal(c.cpu().numpy().view(np.uint32), co)
    # a 2 x 2 x 2 soup scene of device-generated domains against the oracle's own copy of it
    sc = scenes.soup_scene_device(grid=(2, 2, 2), tris_per_domain=1 << 15)
    doms_o = []
    for i, d in enumerate(sc["domains"]):
        b = np.asarray(d["bounds"], np.float32)
        vo, fo, co = oracle.gen_soup(0x5EED0000 + i, 1 << 15, b[:3], b[3:], scenes.soup_h(1 << 15, b[:3], b[3:]))
        doms_o.append(dict(verts=vo, faces=fo, rgb=co, transform=None, bounds=b, mtl=d["mtl"]))
    osc = oracle.Scene(doms_o)
    dev = capi.Device(0); dev.load_scene(sc)
    rays = random_rays(100000, (0, 0, 0), (40, 40, 40), seed=21)
    keys = np.zeros(len(rays), np.uint64)
    dev.intersect_domains(rays, len(rays), keys, None)
    ko, _ = osc.intersect(rays)
    assert np.array_equal(keys, ko) and (keys != oracle.KEY_MISS).mean() > 0.5
    dev.close()
