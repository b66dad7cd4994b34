"""GPU tests of the multi-GPU section of the C ABI (pytest -m gpu).

* nranks = 1: spraygpu_trace_tile_insitu (wavefront kernels + NCCL calls on a 1-rank communicator) must
  reproduce the fused persistent kernel — ray counts exactly, image to float-atomic-order accuracy.
* >= 2 GPUs visible: tests/multigpu_check.py under torchrun, 2 ranks (domain sharding + NCCL min-all-reduce
  on the hit keys, image partition + NCCL gather), compared with the single-GPU render."""
import os
import subprocess
import sys

import numpy as np
import pytest

from spray_b200 import capi, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rmse(a, b):
    return float(np.sqrt(np.mean((a[..., :3].astype(np.float64) - b[..., :3].astype(np.float64)) ** 2)))


@pytest.mark.parametrize("p2p", [False, True, "fwd"])
@pytest.mark.parametrize("kw", [dict(bounces=1, pixel_samples=1), dict(bounces=3, pixel_samples=2),
                                dict(shader=1, bounces=2, pixel_samples=1, ao_samples=3)])
def test_insitu_single_rank_equals_fused(kw, p2p):
    import torch
    sc = scenes.wavelets64(16)
    if kw.get("shader", 0) == 0:
        sc["lights"] = sc["lights"] + [(scenes.LIGHT_DIFFUSE, (0, 0, 0), (0.3, 0.3, 0.3))] if kw["bounces"] == 3 else sc["lights"]
    mt = [(scenes.MTL_DIFFUSE, (1, 1, 1)), (scenes.MTL_MIRROR, (1, 1, 1)), (scenes.MTL_TRANSMISSION, (1.0, 1.5, 0))]
    if kw["bounces"] == 3:
        for i, d in enumerate(sc["domains"]):
            d["mtl"] = mt[i % 3]
    W = 160
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cd = scenes.camera_frame(pos + np.array([25, 15, 5], np.float32), lookat, up, 75.0, W, W)
    cam = capi.make_camera(cd)
    # the in-situ tile always jitters with RandomSampler_init(pixid, s): give the fused kernel the same seed
    sh = capi.make_shade(lights=sc["lights"], eye_seed=capi.EYE_SEED_INSITU, **dict(dict(ao_samples=2), **kw), **scenes.WAVELETS64_BLINN)
    dev = capi.Device(0)
    dev.load_scene(sc)
    comm = capi.Comm(dev, 1, 0, capi.comm_unique_id())
    tiles = scenes.image_schedule_tiles(W, W, sh.pixel_samples, 1, 0, 1 << 13)
    fb_a = torch.zeros((W, W, 4), dtype=torch.float32, device="cuda")
    fb_b = torch.zeros_like(fb_a)
    ta, tb = {}, {}
    for t in tiles:
        for fb, tot, fn in ((fb_a, ta, lambda: dev.trace_tile(cam, sh, t, fb_a, stats=True)),
                            (fb_b, tb, (lambda: comm.trace_tile_insitu_fwd(cam, sh, t, fb_b, 8, [0] * 16, stats=True)) if p2p == "fwd" else
                                 (lambda: comm.trace_tile_insitu_p2p(cam, sh, t, fb_b, 8, stats=True)) if p2p else
                                 (lambda: comm.trace_tile_insitu(cam, sh, t, fb_b, stats=True)))):
            for k, v in fn().as_dict().items():
                tot[k] = tot.get(k, 0) + v
    dev.synchronize()
    for k in ("eye_rays", "radiance_rays", "shadow_rays", "hits"):
        assert ta[k] == tb[k], (k, ta[k], tb[k])
    a, b = fb_a.cpu().numpy(), fb_b.cpu().numpy()
    assert a[..., :3].max() > 0.05
    assert _rmse(a, b) <= 1e-6
    comm.image_reduce(fb_b, W * W, 0)          # 1-rank collectives are no-ops
    comm.close(); dev.close()


@pytest.mark.parametrize("spp,bounces", [(4, 1), (2, 3), (1, 2)])
def test_insitu_single_rank_vs_oracle(oracle, spp, bounces):
    """a13 against the ORACLE (VERDICT r1 "What's weak" 2): the in-situ tile at one rank, C5's shape (wavelet domains,
    multi-sample), with the in-situ tracers' eye-ray jitter seed RandomSampler_init(pixid, s)
    (src/insitu/insitu_ray.h:161-165) — NOT the ooc tracer's RandomSampler_init(bufid) the first round used."""
    import torch
    sc = scenes.wavelets64(8)
    W, H = 144, 112
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cd = scenes.camera_frame(pos, lookat, up, 90.0, W, H)
    kw = dict(bounces=bounces, pixel_samples=spp, **scenes.WAVELETS64_BLINN)
    cam_o, cam = oracle.camera_from_desc(cd), capi.make_camera(cd)
    tiles = scenes.image_schedule_tiles(W, H, spp, 1, 0, 1 << 13)
    assert len(tiles) > 1
    # eye rays: bit-exact, and different from the ooc seed at spp > 1
    dev = capi.Device(0)
    dev.load_scene(sc)
    for t in tiles[:2]:
        rays = np.zeros((t[2] * t[3] * spp, 8), np.float32)
        dev.gen_eye_rays(cam, t, spp, rays, insitu=True)
        assert np.array_equal(rays, oracle.gen_eye_rays(cam_o, t, spp, eye_seed=1))
        if spp > 1:
            assert not np.array_equal(rays, oracle.gen_eye_rays(cam_o, t, spp, eye_seed=0))
    sh_o, keep = oracle.make_shade(lights=sc["lights"], eye_seed=1, **kw)
    fb_o, st_o = oracle.Scene(sc["domains"]).render(cam_o, sh_o, tiles)
    comm = capi.Comm(dev, 1, 0, capi.comm_unique_id())
    sh = capi.make_shade(lights=sc["lights"], **kw)          # trace_tile_insitu forces the in-situ seed
    fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    tot = {}
    for t in tiles:
        for k, v in comm.trace_tile_insitu(cam, sh, t, fb, stats=True).as_dict().items():
            tot[k] = tot.get(k, 0) + v
    dev.synchronize()
    assert tot["eye_rays"] == st_o.eye_rays == W * H * spp
    assert tot["radiance_rays"] == st_o.radiance_rays and tot["shadow_rays"] == st_o.shadow_rays
    assert fb_o[..., :3].max() > 0.1
    assert _rmse(fb.cpu().numpy(), fb_o) <= 1e-4
    # the peer-memory variant (insitu2.cu): same image and counts
    fb2 = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    tot2 = {}
    for t in tiles:
        for k, v in comm.trace_tile_insitu_p2p(cam, sh, t, fb2, 4, stats=True).as_dict().items():
            tot2[k] = tot2.get(k, 0) + v
    dev.synchronize()
    assert tot2["radiance_rays"] == st_o.radiance_rays and tot2["shadow_rays"] == st_o.shadow_rays and tot2["hits"] == tot["hits"]
    assert _rmse(fb2.cpu().numpy(), fb_o) <= 1e-4
    # the ray-forwarding variant (insitu3.cu): same image and counts
    fb3 = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    tot3 = {}
    for t in tiles:
        for k, v in comm.trace_tile_insitu_fwd(cam, sh, t, fb3, 4, [0] * len(sc["domains"]), stats=True).as_dict().items():
            tot3[k] = tot3.get(k, 0) + v
    dev.synchronize()
    assert tot3["radiance_rays"] == st_o.radiance_rays and tot3["shadow_rays"] == st_o.shadow_rays and tot3["hits"] == tot["hits"]
    assert _rmse(fb3.cpu().numpy(), fb_o) <= 1e-4
    # the fused kernel follows shade.eye_seed (bench.py's C5 at one GPU)
    sh1 = capi.make_shade(lights=sc["lights"], eye_seed=capi.EYE_SEED_INSITU, **kw)
    fb1 = np.zeros((H, W, 4), np.float32)
    t1 = {}
    for t in tiles:
        for k, v in dev.trace_tile(cam, sh1, t, fb1, stats=True).as_dict().items():
            t1[k] = t1.get(k, 0) + v
    assert t1["radiance_rays"] == st_o.radiance_rays and t1["shadow_rays"] == st_o.shadow_rays
    assert _rmse(fb1, fb_o) <= 1e-4
    if spp > 1:
        sh_o0, keep0 = oracle.make_shade(lights=sc["lights"], eye_seed=0, **kw)
        fb_o0, _ = oracle.Scene(sc["domains"]).render(cam_o, sh_o0, tiles)
        assert _rmse(fb_o0, fb_o) > 1e-4                        # the two seeds really give different images
    comm.close(); dev.close()


def test_fwd_inbox_overflow_is_reported_once_and_cleared(monkeypatch):
    """Ray forwarding: an inbox overflow (rays dropped) must surface even on the timed path, which passes stats = NULL —
    the tile's last kernel mirrors it into the mapped error word, the NEXT call reports it once (SPRAYGPU_ERR_LIMIT)
    and clears it, and the communicator works again afterwards.  SPRAY_FWD_SEG_CAP shrinks the segments for the test."""
    import torch
    sc = scenes.wavelets64(4)
    dev = capi.Device(0); dev.load_scene(sc)
    comm = capi.Comm(dev, 1, 0, capi.comm_unique_id())
    W = H = 64
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cam = capi.make_camera(scenes.camera_frame(pos, lookat, up, 90.0, W, H))
    sh = capi.make_shade(lights=sc["lights"], bounces=1, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    owner = [0] * len(sc["domains"])
    fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    good = comm.trace_tile_insitu_fwd(cam, sh, (0, 0, W, H), fb, 4, owner, stats=True).as_dict()
    dev.synchronize()
    want = fb.cpu().numpy().copy()
    assert good["hits"] > 64
    monkeypatch.setenv("SPRAY_FWD_SEG_CAP", "16")
    fb.zero_()
    comm.trace_tile_insitu_fwd(cam, sh, (0, 0, W, H), fb, 4, owner)                   # no stats: returns OK, rays were dropped
    dev.synchronize()
    monkeypatch.delenv("SPRAY_FWD_SEG_CAP")
    with pytest.raises(capi.SprayGpuError, match="overflowed"):                        # ... and the next call says so
        comm.trace_tile_insitu_fwd(cam, sh, (0, 0, W, H), fb, 4, owner)
    fb.zero_()
    again = comm.trace_tile_insitu_fwd(cam, sh, (0, 0, W, H), fb, 4, owner, stats=True).as_dict()   # reported once, then clean
    dev.synchronize()
    assert again == good and np.array_equal(fb.cpu().numpy(), want)
    monkeypatch.setenv("SPRAY_FWD_SEG_CAP", "16")
    with pytest.raises(capi.SprayGpuError, match="overflowed"):                        # with stats the same call reports it
        comm.trace_tile_insitu_fwd(cam, sh, (0, 0, W, H), fb, 4, owner, stats=True)
    monkeypatch.delenv("SPRAY_FWD_SEG_CAP")
    fb.zero_()
    assert comm.trace_tile_insitu_fwd(cam, sh, (0, 0, W, H), fb, 4, owner, stats=True).as_dict() == good
    comm.close(); dev.close()


def test_insitu_rejects_glass_and_host_fb():
    sc = scenes.wavelets64(2)
    sc["domains"][1]["mtl"] = (scenes.MTL_GLASS, (1.0, 1.5, 0))
    dev = capi.Device(0); dev.load_scene(sc)
    comm = capi.Comm(dev, 1, 0, capi.comm_unique_id())
    cd = scenes.camera_frame((40, 10, 40), (10, 0, 0), (0, 1, 0), 60.0, 32, 32)
    sh = capi.make_shade(lights=sc["lights"], bounces=2, **scenes.WAVELETS64_BLINN)
    import torch
    fb = torch.zeros((32, 32, 4), dtype=torch.float32, device="cuda")
    with pytest.raises(capi.SprayGpuError):
        comm.trace_tile_insitu(capi.make_camera(cd), sh, (0, 0, 32, 32), fb)
    with pytest.raises(capi.SprayGpuError):
        comm.trace_tile_insitu(capi.make_camera(cd), sh, (0, 0, 32, 32), np.zeros((32, 32, 4), np.float32))
    comm.close(); dev.close()


def test_two_ranks_torchrun():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tests", "multigpu_check.py")],
                       capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-3000:])
    assert r.returncode == 0 and "MULTIGPU_CHECK_OK" in r.stdout
