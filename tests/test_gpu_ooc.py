"""GPU parity of the out-of-core tile (spraygpu_trace_tile_ooc; config C4's
mechanism at test size): a device cache of `capacity` blocks driven by an LRU (reference src/render/lru_cache.cc)
must give the oracle's image and ray counts for every capacity, including 1 (every domain visit is a miss ->
mesh upload + BVH build, the rtcCommit-on-miss of trimesh_buffer.cc:117-225)."""
import numpy as np
import pytest
import torch

from spray_b200 import capi, scenes

pytestmark = pytest.mark.gpu
RMSE_TOL = 1e-4


def _rmse(a, b):
    return float(np.sqrt(np.mean((np.asarray(a, np.float64)[..., :3] - np.asarray(b, np.float64)[..., :3]) ** 2)))


def _render_ooc(sc, cd, kw, tiles, capacity, group=None):
    dev = capi.Device(0)
    cache = capi.LruDomainCache(dev, sc, capacity)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **kw)
    fb = torch.zeros((cd["image_h"], cd["image_w"], 4), dtype=torch.float32, device="cuda")
    tot = {}
    for t in tiles:
        st = dev.trace_tile_ooc(cam, sh, t, fb, group or capacity, cache.load, stats=True)
        for k, v in st.as_dict().items():
            tot[k] = tot.get(k, 0) + v
    dev.synchronize()
    out = fb.cpu().numpy()
    dev.close()
    return out, tot, cache


def _oracle(oracle, sc, cd, kw, tiles):
    osc = oracle.Scene(sc["domains"])
    sh_o, keep = oracle.make_shade(lights=sc["lights"], **kw)
    return osc.render(oracle.camera_from_desc(cd), sh_o, tiles)


@pytest.mark.parametrize("capacity", [1, 5, 16, 64])
@pytest.mark.parametrize("kw", [dict(bounces=1, pixel_samples=1), dict(bounces=3, pixel_samples=2)])
def test_ooc_equals_oracle_for_every_cache_size(oracle, capacity, kw):
    sc = scenes.wavelets64()
    p = scenes.WAVELETS64_CAMERA
    W = 160
    cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, W)
    kw = dict(kw, **scenes.WAVELETS64_BLINN)
    tiles = scenes.image_schedule_tiles(W, W, kw["pixel_samples"], 1, 0, 1 << 14)
    fb_o, st_o = _oracle(oracle, sc, cd, kw, tiles)
    fb_g, st_g, cache = _render_ooc(sc, cd, kw, tiles, capacity)
    assert st_g["eye_rays"] == st_o.eye_rays
    assert st_g["radiance_rays"] == st_o.radiance_rays
    assert st_g["shadow_rays"] == st_o.shadow_rays
    assert 0 < st_g["hits"] <= st_g["radiance_rays"]
    assert _rmse(fb_g, fb_o) <= RMSE_TOL
    if capacity >= 64:
        assert cache.misses <= 64          # every domain built at most once
    else:
        assert cache.misses > capacity     # the sweeps really cycled the cache


def test_ooc_equals_fused_kernel_bitwise_at_one_sample_per_pixel(oracle):
    """Same arithmetic, same per-pixel accumulation order (depth by depth) -> identical images."""
    sc = scenes.wavelets64()
    p = scenes.WAVELETS64_CAMERA
    W = 128
    cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, W)
    kw = dict(bounces=1, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    tiles = [(0, 0, W, W)]
    fb_ooc, st, _ = _render_ooc(sc, cd, kw, tiles, 7)
    dev = capi.Device(0); dev.load_scene(sc)
    fb = np.zeros((W, W, 4), np.float32)
    st_f = dev.trace_tile(capi.make_camera(cd), capi.make_shade(lights=sc["lights"], **kw), tiles[0], fb, stats=True)
    dev.close()
    assert st["shadow_rays"] == st_f.shadow_rays and st["hits"] == st_f.hits
    assert np.array_equal(fb_ooc, fb)


def test_ooc_area_light_ao_and_mirror(oracle):
    sc = scenes.wavelets64(8)
    mt = [(scenes.MTL_DIFFUSE, (1, 1, 1)), (scenes.MTL_MIRROR, (1, 1, 1)), (scenes.MTL_TRANSMISSION, (1.0, 1.5, 0))]
    for i, dmn in enumerate(sc["domains"]):
        dmn["mtl"] = mt[i % 3]
    sc["lights"] = [(scenes.LIGHT_POINT, (0.0, 500.0, 1000.0), (1.0, 1.0, 1.0)), (scenes.LIGHT_DIFFUSE, (0, 0, 0), (0.5, 0.5, 0.5))]
    W = 96
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cd = scenes.camera_frame(pos + np.array([30, 20, 10], np.float32), lookat, up, 70.0, W, W)
    for kw in (dict(bounces=3, pixel_samples=2, ao_samples=3), dict(shader=1, bounces=1, pixel_samples=1, ao_samples=4)):
        kw = dict(kw, **scenes.WAVELETS64_BLINN)
        tiles = scenes.image_schedule_tiles(W, W, kw["pixel_samples"], 1, 0, 1 << 13)
        fb_o, st_o = _oracle(oracle, sc, cd, kw, tiles)
        fb_g, st_g, _ = _render_ooc(sc, cd, kw, tiles, 3)
        assert st_g["radiance_rays"] == st_o.radiance_rays and st_g["shadow_rays"] == st_o.shadow_rays
        assert _rmse(fb_g, fb_o) <= RMSE_TOL


def test_ooc_rejects_bad_arguments_and_failing_loader():
    sc = scenes.wavelets64(8)
    dev = capi.Device(0)
    cache = capi.LruDomainCache(dev, sc, 2)
    W = 32
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cd = scenes.camera_frame(pos, lookat, up, 90.0, W, W)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], bounces=1, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    fb = torch.zeros((W, W, 4), dtype=torch.float32, device="cuda")
    with pytest.raises(capi.SprayGpuError):          # host frame buffer
        dev.trace_tile_ooc(cam, sh, (0, 0, W, W), np.zeros((W, W, 4), np.float32), 2, cache.load)
    with pytest.raises(capi.SprayGpuError):          # group larger than the cache: a group member gets evicted
        dev.trace_tile_ooc(cam, sh, (0, 0, W, W), fb, 4, cache.load)
    with pytest.raises(capi.SprayGpuError):          # loader reports failure
        dev.trace_tile_ooc(cam, sh, (0, 0, W, W), fb, 2, lambda i: 3)
    with pytest.raises(ZeroDivisionError):           # loader raises: surfaces in Python, not through the C frame
        dev.trace_tile_ooc(cam, sh, (0, 0, W, W), fb, 2, lambda i: 1 // 0)
    dev.close()


def test_c4_shape_512_domains_fused_and_out_of_core(oracle):
    """Config C4's SHAPE at test size: 8 x 8 x 8 = 512 domains (the limit of the key's 9 domain bits; the top-level tree
    and the domain table need the > 48 KB shared-memory opt-in), 192 triangles each.  The fused kernel with everything
    resident, and the out-of-core sweep behind a 16-block LRU (groups of 8, the next group prefetched while the current
    one is traced; re-loaded blocks replay their build graph), against the oracle: ray counts equal, RMSE <= 1e-4,
    the two GPU paths bit-identical at 1 spp."""
    sc = scenes.soup_scene((8, 8, 8), 192)
    assert len(sc["domains"]) == 512
    W, H = 192, 128
    e = 160.0
    cd = scenes.camera_frame((1.19 * e, 1.12 * e, 1.06 * e), (0.5 * e, 0.5 * e, 0.5 * e), (0, 1, 0), 90.0, W, H)
    kw = dict(bounces=2, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    tiles = scenes.image_schedule_tiles(W, H, 1, 1, 0)
    fb_o, st_o = _oracle(oracle, sc, cd, kw, tiles)
    assert fb_o[..., :3].max() > 0.05 and st_o.shadow_rays > 1000
    # resident, fused
    dev = capi.Device(0)
    dev.load_scene(sc)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **kw)
    fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    st = dev.trace_frame(cam, sh, tiles, fb, stats=True)
    dev.synchronize()
    fused = fb.cpu().numpy()
    dev.close()
    assert st.radiance_rays == st_o.radiance_rays and st.shadow_rays == st_o.shadow_rays
    assert _rmse(fused, fb_o) <= RMSE_TOL
    # out of core: 16 of 512 domains on the device
    img, tot, cache = _render_ooc(sc, cd, kw, tiles, 16)
    assert tot["radiance_rays"] == st_o.radiance_rays and tot["shadow_rays"] == st_o.shadow_rays
    assert cache.misses > 100 and _rmse(img, fb_o) <= RMSE_TOL
    assert _rmse(img, fused) <= 1e-7
