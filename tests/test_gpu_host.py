"""GPU test of the C++ host mirror: host/spray_gpu (SprayRenderer<GpuTracer<Scene<Cache>>>) renders the
reference's wavelets64 example from files on disk, with the reference's command line, and the PPM equals the
oracle's image after the reference's 10-bit quantisation (image.h:183-204)."""
import os
import subprocess

import numpy as np
import pytest

from spray_b200 import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def spray_gpu():
    env = dict(os.environ); env.pop("CC", None); env.pop("CXX", None)
    from spray_b200 import capi
    capi.build()
    subprocess.run(["make", "-C", HOST], check=True, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return os.path.join(HOST, "spray_gpu")


def read_ppm(path):
    toks = open(path).read().split()
    assert toks[0] == "P3" and toks[3] == "1023"
    w, h = int(toks[1]), int(toks[2])
    return np.array(toks[4:], np.int64).reshape(h, w, 3)[::-1]      # rows are written bottom-up


@pytest.mark.parametrize("cache_size,extra", [(-1, []), (64, []), (-1, ["--bounces", "2", "--pixel-samples", "2"]),
                                              (6, []), (3, ["--bounces", "2", "--pixel-samples", "2"])])   # out-of-core LRU
def test_spray_gpu_film_wavelets64(spray_gpu, oracle, tmp_path, cache_size, extra):
    sc = scenes.wavelets64()
    verts, faces, rgb = scenes.load_wavelet_mesh()
    scenes.write_ply(str(tmp_path / "wavelet.ply"), verts, faces, rgb)
    scenes.write_spray_file(sc, str(tmp_path / "w64.spray"))
    W = 256
    p = scenes.WAVELETS64_CAMERA
    out = str(tmp_path / "o.ppm")
    kw = dict(bounces=1, pixel_samples=1)
    for k, v in zip(extra[::2], extra[1::2]):
        kw[k[2:].replace("-", "_")] = int(v)
    cmd = [spray_gpu, "--mode", "film", "-w", str(W), "-h", str(W), "--frames", "2", "--cache-size", str(cache_size),
           "--partition", "image", "--camera"] + [repr(float(x)) for x in p["pos"] + p["lookat"]] + \
          ["--pixel-samples", str(kw["pixel_samples"]), "--bounces", str(kw["bounces"]), "--blinn", "0.4", "0.4", "0.4", "10",
           "--ply-path", str(tmp_path), "-o", out, str(tmp_path / "w64.spray")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "[STATS] total" in r.stdout
    img = read_ppm(out)
    O = oracle
    osc = O.Scene(sc["domains"])
    cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, W)
    sh, keep = O.make_shade(lights=sc["lights"], **kw, **scenes.WAVELETS64_BLINN)
    fb, st = osc.render(O.camera_from_desc(cd), sh, [(0, 0, W, W)])
    ref = np.clip(fb[..., :3].astype(np.float32) * np.float32(1023.0), 0, 1023).astype(np.int64)
    diff = np.abs(img - ref)
    assert diff.max() <= 1                       # float atomics order / powf ulps can flip a quantisation step
    assert (diff > 0).mean() < 1e-3
    assert img.max() > 100                       # not a black frame


def test_spray_gpu_rejects_what_the_reference_rejects(spray_gpu, tmp_path):
    r = subprocess.run([spray_gpu, "--mode", "film", str(tmp_path / "scene.txt")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode != 0 and ".spray" in r.stdout
    r = subprocess.run([spray_gpu, "--mode", "glfw", str(tmp_path / "scene.spray")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode != 0 and "OpenGL" in r.stdout
