"""The streaming variant of the fused tile kernel (SPRAYGPU_STREAM=1, render_tile_stream_kernel) must give the
same rays and image as the oracle.  The switch is read once per process, so the render runs in a subprocess."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, %r)
from oracle import pyoracle as O
from spray_b200 import capi, scenes
sc = scenes.wavelets64()
p = scenes.WAVELETS64_CAMERA
W = 200
cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, W)
osc = O.Scene(sc["domains"])
dev = capi.Device(0); dev.load_scene(sc)
for kw in (dict(bounces=1, pixel_samples=1), dict(bounces=3, pixel_samples=1), dict(bounces=2, pixel_samples=3)):
    kw = dict(kw, **scenes.WAVELETS64_BLINN)
    tiles = scenes.image_schedule_tiles(W, W, kw["pixel_samples"], 1, 0, 1 << 14)
    fb = np.zeros((W, W, 4), np.float32)
    tot = {}
    for t in tiles:
        st = dev.trace_tile(capi.make_camera(cd), capi.make_shade(lights=sc["lights"], **kw), t, fb, stats=True)
        for k, v in st.as_dict().items():
            tot[k] = tot.get(k, 0) + v
    sh, keep = O.make_shade(lights=sc["lights"], **kw)
    fb_o, st_o = osc.render(O.camera_from_desc(cd), sh, tiles)
    rmse = float(np.sqrt(np.mean((fb[..., :3].astype(np.float64) - fb_o[..., :3]) ** 2)))
    assert tot["eye_rays"] == st_o.eye_rays and tot["radiance_rays"] == st_o.radiance_rays and tot["shadow_rays"] == st_o.shadow_rays, (tot, st_o.radiance_rays, st_o.shadow_rays)
    assert rmse <= 1e-4, rmse
    assert fb[..., :3].max() > 0.1
print("STREAM-OK")
""" % ROOT


def test_stream_kernel_equals_oracle():
    env = dict(os.environ, SPRAYGPU_STREAM="1")
    r = subprocess.run([sys.executable, "-c", SCRIPT], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "STREAM-OK" in r.stdout, r.stdout[-3000:]
