"""GPU tests of the frame image (spraygpu_image, image.cu): the device render target + sparse flush into the pinned
host image that replaces HdrImage::buf / clear() / composite() (reference src/display/image.h:58-239).  Single GPU
here; the peer (NVLink) destination is exercised by tests/multigpu_check.py on >= 2 GPUs."""
import numpy as np
import pytest

from spray_b200 import capi, scenes

pytestmark = pytest.mark.gpu


def _frame(dev, sc, cam_desc, kw, fb):
    cam = capi.make_camera(cam_desc)
    sh = capi.make_shade(lights=sc["lights"], **kw)
    tiles = scenes.image_schedule_tiles(cam_desc["image_w"], cam_desc["image_h"], kw["pixel_samples"], 1, 0, 1 << 14)
    dev.trace_frame(cam, sh, tiles, fb)
    return cam, sh, tiles


def test_image_flush_to_host_is_exact_and_sparse():
    """Frames from DIFFERENT cameras: pixels that turn black must be zeroed on the host although the flush only
    stores pixels that are (or were) non-zero; the render target is left cleared; the byte count is what crossed."""
    sc = scenes.wavelets64(8)
    dev = capi.Device(0)
    dev.load_scene(sc)
    W, H = 256, 160
    img = capi.Image(dev, None, W, H)
    kw = dict(bounces=1, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    assert not img.host.any()
    prev_pix = 0
    for k, shift in enumerate(([0, 0, 0], [25, 10, 0], [-40, -5, 30], [0, 0, 0], [300, 300, 300])):
        cpos = pos + np.array(shift, np.float32)
        cd = scenes.camera_frame(cpos, lookat if k < 4 else cpos + (cpos - lookat), up, 90.0, W, H)   # the last one looks away
        _frame(dev, sc, cd, kw, img.device)
        want = img.read_device("device")
        img.flush(to_host=True, clear=True)
        dev.synchronize()
        assert np.array_equal(img.host, want), k
        assert not img.read_device("device").any()                      # image.clear() happened
        nz = int((want != 0).any(axis=-1).sum())
        _, pix = img.flushed_pixels()
        assert nz <= pix - prev_pix <= W * H
        if k == 0:
            assert 0 < pix < W * H // 2                                 # sparse: sky pixels never cross PCIe
        prev_pix = pix
    assert not img.host.any()                                           # the last camera looks away from the scene
    # set_rows: only the listed blocks are flushed
    cd = scenes.camera_frame(pos, lookat, up, 90.0, W, H)
    _frame(dev, sc, cd, kw, img.device)
    want = img.read_device("device")
    img.set_rows([0, 32, 64, 96, 128, 160], [0, -1, 0, -1, 0])          # blocks 1 and 3 belong to nobody here (-1)
    img.flush(to_host=True, clear=True)
    dev.synchronize()
    mine = np.zeros(H, bool); mine[0:32] = mine[64:96] = mine[128:160] = True
    assert np.array_equal(img.host[mine], want[mine]) and not img.host[~mine].any()
    left = img.read_device("device")
    assert not left[mine].any() and np.array_equal(left[~mine], want[~mine])
    with pytest.raises(capi.SprayGpuError):
        capi.Image(dev, None, 100, 50)                                  # width must be a multiple of 32
    img.close(); dev.close()


def test_image_flush_async_pipelines_frames_exactly():
    """spraygpu_image_flush_async: frame k's host-bound pixels are compacted on the device's stream (target cleared) and
    streamed into the host image by a few CTAs on the image's own stream while frame k + 1 is traced.  Frames from
    different cameras, several rounds through both lists: after spraygpu_image_wait the host image is exactly the last
    frame, and the target is left cleared."""
    sc = scenes.wavelets64(8)
    dev = capi.Device(0)
    dev.load_scene(sc)
    W, H = 256, 160
    kw = dict(bounces=1, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    shifts = ([0, 0, 0], [25, 10, 0], [-40, -5, 30], [10, -20, 5], [0, 0, 0], [-15, 25, -10], [5, 5, 5])
    # references: each frame on its own, synchronously
    ref_img = capi.Image(dev, None, W, H)
    want = []
    for sft in shifts:
        cd = scenes.camera_frame(pos + np.array(sft, np.float32), lookat, up, 90.0, W, H)
        _frame(dev, sc, cd, kw, ref_img.device)
        want.append(ref_img.read_device("device"))
        ref_img.flush(clear=True)
    assert not np.array_equal(want[0], want[1])
    img = capi.Image(dev, None, W, H)
    for k, sft in enumerate(shifts):
        cd = scenes.camera_frame(pos + np.array(sft, np.float32), lookat, up, 90.0, W, H)
        _frame(dev, sc, cd, kw, img.device)
        img.flush_async(to_host=True)                # returns at once; the stores overlap the next frame
        if k in (2, 5, 6):                           # look at some frames, let the others stream through
            img.wait()
            assert np.array_equal(img.host, want[k]), k
    img.wait(); dev.synchronize()
    assert not img.read_device("device").any()       # cleared by the compaction pass
    # mixing in a synchronous flush afterwards still works on the current target
    cd = scenes.camera_frame(pos, lookat, up, 90.0, W, H)
    _frame(dev, sc, cd, kw, img.device)
    img.flush(to_host=True, clear=True)
    dev.synchronize()
    assert np.array_equal(img.host, want[0])
    img.close(); ref_img.close(); dev.close()


def test_trace_frame_rects_equals_whole_tiles(oracle):
    """Rectangles of blocking tiles keep their blocking tile's sample ids: any cut of the frame into rectangles gives the
    image of the uncut frame (bit-identical at 1 spp / 1 bounce, same ray counts and RMSE <= 1e-4 vs the oracle at
    4 spp / 3 bounces where float atomics reorder)."""
    import torch
    sc = scenes.wavelets64(16)
    dev = capi.Device(0)
    dev.load_scene(sc)
    W, H = 224, 160
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cd = scenes.camera_frame(pos, lookat, up, 90.0, W, H)
    cam = capi.make_camera(cd)
    for kw in (dict(bounces=1, pixel_samples=1), dict(bounces=3, pixel_samples=4)):
        kw = dict(kw, **scenes.WAVELETS64_BLINN)
        sh = capi.make_shade(lights=sc["lights"], **kw)
        blocking = scenes.image_schedule_tiles(W, H, kw["pixel_samples"], 1, 0, 1 << 13)
        assert len(blocking) > 2
        fb0 = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
        st0 = dev.trace_frame(cam, sh, blocking, fb0, stats=True).as_dict()
        rects, parents = [], []
        for (tx, ty, tw, th) in blocking:                               # ragged cut: rows of 7, columns of 96
            for y in range(ty, ty + th, 7):
                for x in range(0, W, 96):
                    rects.append((x, y, min(96, W - x), min(7, ty + th - y)))
                    parents.append((tx, ty, tw, th))
        assert len(rects) > 32                                          # several launches
        fb1 = torch.zeros_like(fb0)
        st1 = dev.trace_frame_rects(cam, sh, rects, parents, fb1, stats=True).as_dict()
        dev.synchronize()
        for k in ("eye_rays", "radiance_rays", "shadow_rays", "hits"):
            assert st0[k] == st1[k], k
        if kw["pixel_samples"] == 1 and kw["bounces"] == 1:
            assert torch.equal(fb0, fb1)
        sh_o, keep = oracle.make_shade(lights=sc["lights"], **kw)
        fb_o, st_o = oracle.Scene(sc["domains"]).render(oracle.camera_from_desc(cd), sh_o, blocking)
        e = float(np.sqrt(np.mean((fb1.cpu().numpy()[..., :3].astype(np.float64) - fb_o[..., :3]) ** 2)))
        assert e <= 1e-4 and st1["radiance_rays"] == st_o.radiance_rays and st1["shadow_rays"] == st_o.shadow_rays
    with pytest.raises(capi.SprayGpuError):
        dev.trace_frame_rects(cam, sh, [(0, 0, W, 8)], [(0, 4, W, 8)], fb1)     # rectangle outside its blocking tile
    dev.close()


def test_set_rows_releases_rows_after_a_non_black_frame():
    """Re-partitioning AFTER pixels have been flushed: the bitmap of "what the destination holds" is per rank, so rows a
    rank gives up must be zeroed at the destination by that rank (nobody else knows they are non-zero); rows it takes
    back later are flushed from scratch.  Synchronous and asynchronous flushes."""
    sc = scenes.wavelets64(8)
    dev = capi.Device(0)
    dev.load_scene(sc)
    W, H = 256, 160
    kw = dict(bounces=1, pixel_samples=1, **scenes.WAVELETS64_BLINN)
    pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
    cd = scenes.camera_frame(pos, lookat, up, 90.0, W, H)
    for use_async in (False, True):
        img = capi.Image(dev, None, W, H)

        def flush():
            if use_async:
                img.flush_async(to_host=True)        # set_rows itself must wait for the pending stores
            else:
                img.flush(to_host=True, clear=True)
                dev.synchronize()

        _frame(dev, sc, cd, kw, img.device)
        want = img.read_device("device")
        assert want.any()
        flush()
        mine = np.zeros(H, bool); mine[0:32] = mine[64:96] = mine[128:160] = True
        assert want[~mine].any()                                            # the rows given up are not black
        img.set_rows([0, 32, 64, 96, 128, 160], [0, -1, 0, -1, 0])          # blocks 1 and 3 leave this rank
        assert np.array_equal(img.host[mine], want[mine]) and not img.host[~mine].any()
        # a different frame under the new ownership: only my rows arrive, the others stay black
        cd2 = scenes.camera_frame(pos + np.array([25, 10, 0], np.float32), lookat, up, 90.0, W, H)
        _frame(dev, sc, cd2, kw, img.device)
        want2 = img.read_device("device")
        flush(); img.wait()
        assert np.array_equal(img.host[mine], want2[mine]) and not img.host[~mine].any()
        # the rows come back: the leftovers of frame 2 in the target are flushed, black pixels included
        img.set_rows([0, H], None)
        flush(); img.wait()
        assert np.array_equal(img.host[~mine], want2[~mine])
        _frame(dev, sc, cd, kw, img.device)
        flush(); img.wait()
        assert np.array_equal(img.host, want)
        img.close()
    dev.close()
