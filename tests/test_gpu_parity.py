"""GPU parity (run on the B200 box: pytest -m gpu).  Everything goes through the C ABI
(spray_b200.capi -> libspraygpu.so) and is compared with the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): hit (domain, primID) bit-exact against the FP32 oracle O2 (same FP32
contract), >= 95 % exact against the double-precision truth O1 with every mismatch inside O1's
documented near-tie / edge-graze set, |t - t_O1| <= 1e-5 t, image RMSE <= 1e-4."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_rays, random_rays
from spray_b200 import capi, scenes

pytestmark = pytest.mark.gpu
T_REL_TOL = 1e-5      # north_star: hit t within 1e-5 relative
RMSE_TOL = 1e-4       # north_star: image RMSE <= 1e-4


@pytest.fixture(scope="module")
def dev():
    d = capi.Device(0)
    yield d
    d.close()


@pytest.fixture(scope="module")
def c1(oracle, dev):
    sc = scenes.wavelet_single()
    return sc, oracle.Scene(sc["domains"])


@pytest.fixture(scope="module")
def c2(oracle):
    sc = scenes.wavelets64()
    return sc, oracle.Scene(sc["domains"])


def _cam_desc(sc, w, h, pose=None):
    if pose is None:
        pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    else:
        pos, lookat, up = pose
    return scenes.camera_frame(pos, lookat, up, 90.0, w, h)


def _gpu_intersect_domains(dev, rays):
    n = len(rays)
    keys = np.zeros(n, np.uint64); uv = np.zeros((n, 2), np.float32)
    dev.intersect_domains(rays, n, keys, uv)
    return keys, uv


def test_commit_normals_vertices_bit_exact(oracle, dev, c2):
    """a4/a5: transform (glm order) + computeNormals (face-order sums) are bit-identical."""
    sc, osc = c2
    for i in (0, 1, 21, 63):
        d = sc["domains"][i]
        s = dev.new_scene()
        s.set_mesh(d["verts"], d["faces"], d["rgb"]); s.set_transform(d["transform"]); s.commit()
        assert np.array_equal(s.vertices(), osc.verts(i))
        assert np.array_equal(s.normals(), osc.normals(i))
        inf = s.info()
        assert inf.num_faces == 5480 and inf.num_vertices == 2840 and 0 < inf.num_nodes < 5480 // 3
        v = osc.verts(i)
        assert np.array_equal(np.array(list(inf.bounds), np.float32), np.concatenate([v.min(0), v.max(0)]))
        s.close()


def test_single_domain_intersect_occluded(oracle, dev, c1):
    """a1/a2 through seam 2: batched Scene::intersect / Scene::occluded on one domain."""
    sc, osc = c1
    d = sc["domains"][0]
    s = dev.new_scene(); s.set_mesh(d["verts"], d["faces"], d["rgb"]); s.commit()
    cam = oracle.camera_from_desc(_cam_desc(sc, 256, 256))
    rng = np.random.default_rng(7)
    n = 20000
    o = np.round(rng.uniform(-14, 14, (n, 3))).astype(np.float32)
    dd = np.zeros((n, 3), np.float32); dd[np.arange(n), rng.integers(0, 3, n)] = np.where(rng.random(n) < 0.5, 1, -1)
    rays = np.concatenate([oracle.gen_eye_rays(cam, (0, 0, 256, 256)), random_rays(50000, [-10] * 3, [10] * 3, seed=2),
                           make_rays(o, dd)])
    n = len(rays)
    keys = np.zeros(n, np.uint64); uv = np.zeros((n, 2), np.float32); occ = np.zeros(n, np.uint8)
    s.intersect(5, rays, n, keys, uv)      # domain id is only stamped into the key
    s.occluded(rays, n, occ)
    ko, uvo = osc.intersect_domain(0, rays)
    ko = np.where(ko == oracle.KEY_MISS, ko, ko | np.uint64(5 << 23))
    assert np.array_equal(keys, ko)
    assert np.array_equal(uv, uvo)
    assert np.array_equal(occ, osc.occluded_domain(0, rays))
    # ray tmin is honoured, tmax ignored (tfar = inf at every reference call site)
    r2 = rays[:1000].copy(); r2[:, 3] = 15.0; r2[:, 7] = 0.5
    k2 = np.zeros(1000, np.uint64); s.intersect(0, r2, 1000, k2)
    assert np.array_equal(k2, osc.intersect_domain(0, r2)[0])
    # empty batch
    s.intersect(0, rays, 0, keys, uv)
    s.close()


def test_tiny_and_degenerate_meshes(oracle, dev):
    for ntris in (1, 2, 3, 4, 9, 100):
        v, f, rgb = scenes.soup_domain(3, ntris, (0, 0, 0), (20, 20, 20))
        osc = oracle.Scene([dict(verts=v, faces=f, rgb=rgb, transform=None, bounds=np.array([0, 0, 0, 20, 20, 20], np.float32))])
        s = dev.new_scene(); s.set_mesh(v, f, rgb); s.commit()
        rays = random_rays(8000, [0] * 3, [20] * 3, seed=ntris)
        if ntris <= 4:
            tgt = v[np.random.default_rng(1).integers(0, len(v), len(rays))]
            dd = tgt - rays[:, 0:3]; dd /= np.linalg.norm(dd, axis=1, keepdims=True); rays[:, 4:7] = dd
        keys = np.zeros(len(rays), np.uint64); uv = np.zeros((len(rays), 2), np.float32)
        s.intersect(0, rays, len(rays), keys, uv)
        ko, uvo = osc.intersect_domain(0, rays)
        assert np.array_equal(keys, ko) and np.array_equal(uv, uvo), ntris
        s.close()
    # duplicates + zero-area triangles
    base, f0, _ = scenes.soup_domain(9, 300, (0, 0, 0), (10, 10, 10))
    v = np.concatenate([base, base[:300], np.zeros((3, 3), np.float32) + 5.0]).astype(np.float32)
    f = np.concatenate([f0, f0[:100], np.arange(300, dtype=np.uint32).reshape(100, 3) + np.uint32(len(base)),
                        np.array([[len(v) - 3, len(v) - 2, len(v) - 1]], np.uint32)]).astype(np.uint32)
    osc = oracle.Scene([dict(verts=v, faces=f, rgb=None, transform=None, bounds=np.array([0, 0, 0, 10, 10, 10], np.float32))])
    s = dev.new_scene(); s.set_mesh(v, f, None); s.commit()
    rays = random_rays(20000, [0] * 3, [10] * 3, seed=5)
    keys = np.zeros(len(rays), np.uint64); s.intersect(0, rays, len(rays), keys)
    assert np.array_equal(keys, osc.intersect_domain(0, rays)[0])
    assert np.array_equal(s.normals(), osc.normals(0))
    s.close()
    # error behaviour: bad arguments are reported, not crashed on
    s = dev.new_scene()
    with pytest.raises(capi.SprayGpuError):
        s.commit()
    with pytest.raises(capi.SprayGpuError):
        capi._check(capi.lib().spraygpu_scene_set_mesh(s.h, None, 0, None, 0, None))
    s.close()


@pytest.fixture(scope="module")
def dev_c2(c2):
    d = capi.Device(0)
    d.load_scene(c2[0])
    yield d
    d.close()


def test_domain_lists(oracle, dev_c2, c2):
    """a6: all entered domains, (tmin, id) order, entry distances bit-exact (aabb.h:139-169)."""
    sc, osc = c2
    p = scenes.WAVELETS64_CAMERA
    cam = oracle.camera_from_desc(_cam_desc(sc, 200, 200, (p["pos"], p["lookat"], p["up"])))
    b = scenes.scene_bounds(sc)
    inside = random_rays(20000, b[:3], b[3:], seed=3, spread=-0.5)     # origins inside the grid: negative tmin
    rays = np.concatenate([oracle.gen_eye_rays(cam, (0, 0, 200, 200)), random_rays(20000, b[:3], b[3:], seed=1), inside])
    n, cap = len(rays), 24
    counts = np.zeros(n, np.int32); ids = np.full((n, cap), -1, np.int32); tm = np.zeros((n, cap), np.float32)
    dev_c2.domain_lists(rays, n, cap, counts, ids, tm)
    co, io, to = osc.domain_lists(rays, cap)
    assert np.array_equal(counts, co)
    mask = np.arange(cap)[None, :] < counts[:, None]
    assert np.array_equal(ids[mask], io[mask]) and np.array_equal(tm[mask], to[mask])
    assert (tm[mask] < 0).any()


def test_c2_closest_hit_parity(oracle, dev_c2, c2):
    """a1+a6+a7+a12: closest hit over all entered domains == FP32 oracle bit for bit; vs f64 truth."""
    sc, osc = c2
    p = scenes.WAVELETS64_CAMERA
    W = 320
    cam = oracle.camera_from_desc(_cam_desc(sc, W, W, (p["pos"], p["lookat"], p["up"])))
    b = scenes.scene_bounds(sc)
    rays = np.concatenate([oracle.gen_eye_rays(cam, (0, 0, W, W)), random_rays(40000, b[:3], b[3:], seed=11, spread=0.5)])
    keys, uv = _gpu_intersect_domains(dev_c2, rays)
    ko, uvo = osc.intersect(rays)
    assert np.array_equal(keys, ko)
    assert np.array_equal(uv, uvo)
    dom, prim, t, fl = osc.intersect_f64(rays)
    tg, dg, pg = oracle.key_fields(keys)
    hit = keys != oracle.KEY_MISS
    both = hit & (dom >= 0)
    exact = (dg == dom) & (pg == prim)
    assert exact[both].mean() >= 0.95
    assert exact[both & (fl == 0)].all()                      # every mismatch is a documented near-tie/graze ray
    assert (np.abs(tg[both] - t[both]) <= T_REL_TOL * t[both]).all()
    assert (hit != (dom >= 0)).mean() < 2e-3                   # silhouette grazes only
    # occlusion
    occ = np.zeros(len(rays), np.uint8)
    dev_c2.occluded_domains(rays, len(rays), occ)
    assert np.array_equal(occ, osc.occluded(rays))
    # hit attributes (a3): colour truncation + unnormalised Ns, (u,v) re-derived from the key
    color = np.zeros(len(rays), np.uint32); ns = np.zeros((len(rays), 3), np.float32)
    dev_c2.hit_attributes(rays, keys, len(rays), color, ns)
    co, no = osc.hit_attributes(ko, uvo)
    assert np.array_equal(color, co) and np.array_equal(ns, no)


def test_merge_mode_equals_single_pass(oracle, c2):
    """Speculative resolution: tracing two disjoint domain subsets into the same key buffer with the 64-bit
    atomicMin merge gives the single-pass result (what the NCCL min-all-reduce does across GPUs)."""
    sc, osc = c2
    p = scenes.WAVELETS64_CAMERA
    cam = oracle.camera_from_desc(_cam_desc(sc, 192, 192, (p["pos"], p["lookat"], p["up"])))
    rays = oracle.gen_eye_rays(cam, (0, 0, 192, 192))
    n = len(rays)
    ko, _ = osc.intersect(rays)
    occ_o = osc.occluded(rays)
    keys = np.full(n, oracle.KEY_MISS, np.uint64); occ = np.zeros(n, np.uint8)
    for part in (range(0, 64, 2), range(1, 64, 2)):
        d = capi.Device(0)
        d.load_scene(sc, owned=part)
        d.intersect_domains(rays, n, keys, None, merge=True)
        d.occluded_domains(rays, n, occ, merge=True)
        d.close()
    assert np.array_equal(keys, ko)
    assert np.array_equal(occ, occ_o)


def test_eye_rays(oracle, dev, c1):
    """a9: pixel-corner rays (spp 1) and jittered rays (spp 4, seed = bufid) are bit-identical."""
    sc, _ = c1
    cd = _cam_desc(sc, 96, 64)
    cam_o, cam_g = oracle.camera_from_desc(cd), capi.make_camera(cd)
    for spp, tile in ((1, (0, 0, 96, 64)), (4, (8, 16, 40, 24)), (2, (95, 63, 1, 1))):
        rays = np.zeros((tile[2] * tile[3] * spp, 8), np.float32)
        dev.gen_eye_rays(cam_g, tile, spp, rays)
        assert np.array_equal(rays, oracle.gen_eye_rays(cam_o, tile, spp))


def _render_pair(oracle, dev, osc, sc, cd, shade_kw, tiles, host_fb=True):
    cam_o, cam_g = oracle.camera_from_desc(cd), capi.make_camera(cd)
    sh_o, keep = oracle.make_shade(lights=sc["lights"], **shade_kw)
    sh_g = capi.make_shade(lights=sc["lights"], **shade_kw)
    fb_o, st_o = osc.render(cam_o, sh_o, tiles)
    fb_g = np.zeros_like(fb_o)
    tot = {}
    for t in tiles:
        st = dev.trace_tile(cam_g, sh_g, t, fb_g, stats=True)
        for k, v in st.as_dict().items():
            tot[k] = tot.get(k, 0) + v
    return fb_o, st_o, fb_g, tot


def _rmse(a, b):
    return float(np.sqrt(np.mean((a[..., :3].astype(np.float64) - b[..., :3].astype(np.float64)) ** 2)))


def test_render_c1(oracle, c1):
    """Config C1 shape: wavelet.ply, default camera, PT shader, 1 point light, bounces 1, blinn 0.4/10."""
    sc, osc = c1
    d = capi.Device(0); d.load_scene(sc)
    W = 256
    cd = _cam_desc(sc, W, W)
    tiles = scenes.image_schedule_tiles(W, W, 1, 1, 0, 1 << 14)        # 4 blocking tiles
    fb_o, st_o, fb_g, st_g = _render_pair(oracle, d, osc, sc, cd, dict(bounces=1, pixel_samples=1, **scenes.WAVELETS64_BLINN), tiles)
    assert st_g["eye_rays"] == st_o.eye_rays == W * W
    assert st_g["radiance_rays"] == st_o.radiance_rays and st_g["shadow_rays"] == st_o.shadow_rays
    assert _rmse(fb_g, fb_o) <= RMSE_TOL
    assert np.abs(fb_g - fb_o).max() <= 2e-5        # only powf differs (libm vs CUDA), a few ulp
    d.close()


@pytest.mark.parametrize("kw,W", [(dict(bounces=1, pixel_samples=1), 192), (dict(bounces=3, pixel_samples=1), 192),
                                  (dict(bounces=2, pixel_samples=4), 192),
                                  (dict(shader=1, bounces=1, pixel_samples=1, ao_samples=4), 192),
                                  (dict(shader=1, bounces=2, pixel_samples=2, ao_samples=2), 192),
                                  # the two cases round 2's 2-rank check missed the 1e-4 bar on (1.4e-4 / 2.7e-4 with
                                  # libdevice's cosf / sinf / powf against glibc's), at that check's image size
                                  (dict(bounces=3, pixel_samples=2), 384),
                                  (dict(shader=1, bounces=2, pixel_samples=2, ao_samples=3), 384)])
def test_render_c2(oracle, dev_c2, c2, kw, W):
    """Config C2 shape (wavelets64 camera/blinn): PT and AO shaders, bounces, multi-sampling; RMSE <= 1e-4 (north_star's
    tolerance) and equal ray counts in every case."""
    sc, osc = c2
    p = scenes.WAVELETS64_CAMERA
    cd = _cam_desc(sc, W, W, (p["pos"], p["lookat"], p["up"]))
    spp = kw.get("pixel_samples", 1)
    tiles = scenes.image_schedule_tiles(W, W, spp, 1, 0, 1 << 14)
    fb_o, st_o, fb_g, st_g = _render_pair(oracle, dev_c2, osc, sc, cd, dict(kw, **scenes.WAVELETS64_BLINN), tiles)
    assert st_g["eye_rays"] == st_o.eye_rays == W * W * spp
    assert st_g["radiance_rays"] == st_o.radiance_rays
    assert st_g["shadow_rays"] == st_o.shadow_rays
    assert fb_o[..., :3].max() > 0.1
    assert _rmse(fb_g, fb_o) <= RMSE_TOL


def test_render_area_light_and_materials(oracle):
    """diffuse-hemisphere light (area sampling, RNG-seeded) + mirror / transmission / glass domains."""
    sc = scenes.wavelets64(8)
    mt = [(scenes.MTL_DIFFUSE, (1, 1, 1)), (scenes.MTL_MIRROR, (1, 1, 1)), (scenes.MTL_TRANSMISSION, (1.0, 1.5, 0)),
          (scenes.MTL_GLASS, (1.0, 1.5, 0))]
    for i, dmn in enumerate(sc["domains"]):
        dmn["mtl"] = mt[i % 4]
    sc["lights"] = [(scenes.LIGHT_POINT, (0.0, 500.0, 1000.0), (1.0, 1.0, 1.0)), (scenes.LIGHT_DIFFUSE, (0, 0, 0), (0.5, 0.5, 0.5))]
    osc = oracle.Scene(sc["domains"])
    d = capi.Device(0); d.load_scene(sc)
    W = 128
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cd = scenes.camera_frame(pos + np.array([30, 20, 10], np.float32), lookat, up, 70.0, W, W)
    tiles = scenes.image_schedule_tiles(W, W, 2, 1, 0, 1 << 13)
    kw = dict(bounces=4, pixel_samples=2, ao_samples=3, **scenes.WAVELETS64_BLINN)
    fb_o, st_o, fb_g, st_g = _render_pair(oracle, d, osc, sc, cd, kw, tiles)
    assert st_g["radiance_rays"] == st_o.radiance_rays and st_g["shadow_rays"] == st_o.shadow_rays
    assert st_o.radiance_rays > st_o.eye_rays
    assert _rmse(fb_g, fb_o) <= RMSE_TOL
    d.close()


def test_device_pointers_and_idempotence(oracle, dev_c2, c2):
    """Device-resident buffers (torch CUDA tensors) give the same answers as host buffers; rendering the
    same tile twice into a cleared device frame buffer is bit-identical at 1 spp / 1 light (one add per
    pixel), and accumulates (doubles) without a clear."""
    import torch
    sc, osc = c2
    p = scenes.WAVELETS64_CAMERA
    W = 256
    cd = _cam_desc(sc, W, W, (p["pos"], p["lookat"], p["up"]))
    cam_g = capi.make_camera(cd)
    n = W * W
    rays_d = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
    dev_c2.gen_eye_rays(cam_g, (0, 0, W, W), 1, rays_d)
    keys_d = torch.zeros(n, dtype=torch.int64, device="cuda")
    uv_d = torch.zeros((n, 2), dtype=torch.float32, device="cuda")
    dev_c2.intersect_domains(rays_d, n, keys_d, uv_d)
    dev_c2.synchronize()
    rays_h = rays_d.cpu().numpy()
    kh, uvh = _gpu_intersect_domains(dev_c2, rays_h)
    assert np.array_equal(keys_d.cpu().numpy().view(np.uint64), kh) and np.array_equal(uv_d.cpu().numpy(), uvh)
    sh = capi.make_shade(bounces=1, pixel_samples=1, lights=sc["lights"], **scenes.WAVELETS64_BLINN)
    fb1 = torch.zeros((W, W, 4), dtype=torch.float32, device="cuda")
    dev_c2.trace_tile(cam_g, sh, (0, 0, W, W), fb1)
    fb2 = torch.zeros_like(fb1)
    dev_c2.trace_tile(cam_g, sh, (0, 0, W, W), fb2)
    dev_c2.synchronize()
    assert torch.equal(fb1, fb2)
    dev_c2.trace_tile(cam_g, sh, (0, 0, W, W), fb2)
    dev_c2.synchronize()
    assert torch.equal(fb2, 2 * fb1)
    fb_h = np.zeros((W, W, 4), np.float32)
    dev_c2.trace_tile(cam_g, sh, (0, 0, W, W), fb_h)
    assert np.array_equal(fb_h, fb1.cpu().numpy())
    # tiles partition the image: 7 ragged tiles == one full tile
    fb3 = torch.zeros_like(fb1)
    for t in [(0, 0, 100, 37), (100, 0, 156, 37), (0, 37, 256, 1), (0, 38, 3, 218), (3, 38, 253, 100), (3, 138, 253, 118), (0, 0, 0, 0)]:
        dev_c2.trace_tile(cam_g, sh, t, fb3)
    dev_c2.synchronize()
    assert torch.equal(fb3, fb1)


def test_full_size_properties_c2(oracle, dev_c2, c2):
    """BASELINE config C2 at full size (2048^2): size-independent checks — counters add up, the image is a
    function of the pixel only (tile decomposition invariant), shadowed pixels are exactly black, and a
    stride-16 subsample of the keys equals the oracle."""
    import torch
    sc, osc = c2
    p = scenes.WAVELETS64_CAMERA
    W = 2048
    cd = _cam_desc(sc, W, W, (p["pos"], p["lookat"], p["up"]))
    cam_g = capi.make_camera(cd)
    sh = capi.make_shade(bounces=1, pixel_samples=1, lights=sc["lights"], **scenes.WAVELETS64_BLINN)
    fb = torch.zeros((W, W, 4), dtype=torch.float32, device="cuda")
    tot = {}
    for t in scenes.image_schedule_tiles(W, W, 1, 1, 0):
        st = dev_c2.trace_tile(cam_g, sh, t, fb, stats=True)
        for k, v in st.as_dict().items():
            tot[k] = tot.get(k, 0) + v
    assert tot["eye_rays"] == W * W == tot["radiance_rays"]
    assert tot["shadow_rays"] <= tot["hits"] <= W * W
    n = W * W
    rays_d = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
    dev_c2.gen_eye_rays(cam_g, (0, 0, W, W), 1, rays_d)
    keys_d = torch.zeros(n, dtype=torch.int64, device="cuda")
    dev_c2.intersect_domains(rays_d, n, keys_d, None)
    dev_c2.synchronize()
    keys = keys_d.cpu().numpy().view(np.uint64)
    assert int((keys != oracle.KEY_MISS).sum()) == tot["hits"]
    img = fb.cpu().numpy()
    assert not img[..., :3].reshape(-1, 3)[keys == oracle.KEY_MISS].any()
    sub = np.arange(0, n, 16)
    rays_h = rays_d.cpu().numpy()
    ko, _ = osc.intersect(rays_h[sub])
    assert np.array_equal(keys[sub], ko)
    fb2 = torch.zeros_like(fb)
    dev_c2.trace_tile(cam_g, sh, (0, 0, W, W), fb2)
    dev_c2.synchronize()
    assert torch.equal(fb, fb2)


@pytest.mark.gpu
def test_trace_tile_async_host_frame_buffer(oracle, dev_c2, c2):
    """spraygpu_trace_tile_async (overlapped D2H on a second stream, two staging images): same image as the
    synchronous call, over several frames so that both staging images are reused."""
    import torch
    from spray_b200 import capi
    sc = c2[0]
    p = scenes.WAVELETS64_CAMERA
    W = 192
    cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, W)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(bounces=1, pixel_samples=1, lights=sc["lights"], **scenes.WAVELETS64_BLINN)
    tiles = [(0, y, W, 48) for y in range(0, W, 48)]
    ref = np.zeros((W, W, 4), np.float32)
    for t in tiles:
        dev_c2.trace_tile(cam, sh, t, ref)
    fb = torch.zeros((W, W, 4), dtype=torch.float32).pin_memory()
    fbn = fb.numpy()
    for frame in range(3):
        fbn[:] = -1.0
        for t in tiles:
            dev_c2.trace_tile_async(cam, sh, t, fbn)
        dev_c2.synchronize()
        assert np.abs(fbn[..., :3] - ref[..., :3]).max() <= 1e-6
        assert (fbn[..., 3] == 0).all()
    assert ref[..., :3].max() > 0.1


@pytest.mark.parametrize("kw", [dict(bounces=1, pixel_samples=1), dict(bounces=3, pixel_samples=1), dict(bounces=2, pixel_samples=3)])
def test_trace_frame_equals_tile_loop(oracle, dev_c2, c2, kw):
    """spraygpu_trace_frame (all blocking tiles in one persistent launch) == the per-tile loop: same counters, same
    image bit for bit at 1 spp (one atomic add per pixel and depth), RMSE <= 1e-4 vs the oracle otherwise; ragged
    tile sizes, > 32 tiles (several launches), host frame buffer and the empty frame included."""
    import torch
    sc, osc = c2
    p = scenes.WAVELETS64_CAMERA
    W, H = 200, 150
    cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, H)
    cam_g = capi.make_camera(cd)
    kw = dict(kw, **scenes.WAVELETS64_BLINN)
    spp = kw["pixel_samples"]
    sh = capi.make_shade(lights=sc["lights"], **kw)
    for cap in (1 << 12, 700):                       # 8 tiles ; 43+ tiles of 3 rows (ragged last one)
        tiles = scenes.image_schedule_tiles(W, H, spp, 1, 0, cap)
        assert (len(tiles) > 32) == (cap == 700)
        fb1 = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
        tot = {}
        for t in tiles:
            st = dev_c2.trace_tile(cam_g, sh, t, fb1, stats=True)
            for k, v in st.as_dict().items():
                tot[k] = tot.get(k, 0) + v
        fb2 = torch.zeros_like(fb1)
        st2 = dev_c2.trace_frame(cam_g, sh, tiles, fb2, stats=True).as_dict()
        dev_c2.synchronize()
        for k in ("eye_rays", "radiance_rays", "shadow_rays", "hits"):
            assert st2[k] == tot[k], k
        if spp == 1 and kw["bounces"] == 1:
            assert torch.equal(fb1, fb2)
        sh_o, keep = oracle.make_shade(lights=sc["lights"], **kw)
        fb_o, st_o = osc.render(oracle.camera_from_desc(cd), sh_o, tiles)
        assert st2["radiance_rays"] == st_o.radiance_rays and st2["shadow_rays"] == st_o.shadow_rays
        assert _rmse(fb2.cpu().numpy(), fb_o) <= RMSE_TOL
        fbh = np.zeros((H, W, 4), np.float32)
        dev_c2.trace_frame(cam_g, sh, tiles, fbh)     # host frame buffer path
        assert _rmse(fbh, fb_o) <= RMSE_TOL
    fb3 = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    st3 = dev_c2.trace_frame(cam_g, sh, [], fb3, stats=True).as_dict()
    assert st3["eye_rays"] == 0 and not fb3.any()
