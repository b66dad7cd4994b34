import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure).  Built on demand with gcc."""
    from oracle import pyoracle
    pyoracle.build()
    pyoracle.lib()
    return pyoracle


def _load_hostemu(soname, flags):
    import ctypes as C
    d = os.path.join(ROOT, "tests", "hostemu")
    so = os.path.join(d, soname)
    src = os.path.join(d, "hostemu.cpp")
    core = os.path.join(ROOT, "spray_b200", "csrc", "core.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(core)):
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.run([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-mavx2", "-mfma",
                        "-Wno-unknown-pragmas"] + flags + ["-o", so, src], check=True)
    E = C.CDLL(so)
    E.emu_build.restype = C.c_void_p
    E.emu_build.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    E.emu_destroy.argtypes = [C.c_void_p]
    E.emu_num_nodes.restype = C.c_size_t
    E.emu_num_nodes.argtypes = [C.c_void_p]
    E.emu_get_bvh.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    E.emu_depth.restype = C.c_int
    E.emu_depth.argtypes = [C.c_void_p]
    E.emu_intersect.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    E.emu_occluded.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    E.emu_rng_init1.restype = C.c_uint32
    E.emu_rng_init1.argtypes = [C.c_uint32]
    E.emu_rng_float.restype = C.c_float
    E.emu_rng_float.argtypes = [C.POINTER(C.c_uint32)]
    E.emu_morton.restype = C.c_uint32
    E.emu_morton.argtypes = [C.c_float] * 3
    E.emu_domain_slab.restype = C.c_int
    E.emu_domain_slab.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_float)]
    E.emu_interp_color.restype = C.c_uint32
    E.emu_interp_color.argtypes = [C.c_uint32] * 3 + [C.c_float] * 2
    E.emu_cosine_hemisphere.argtypes = [C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.POINTER(C.c_float)]
    E.emu_blinn_phong.argtypes = [C.c_float, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    E.emu_camera_dir.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p]
    return E


@pytest.fixture(scope="session")
def hostemu():
    """Host emulation of spray_b200/csrc/core.cuh (development harness, tests/hostemu)."""
    return _load_hostemu("libhostemu.so", [])


@pytest.fixture(scope="session")
def hostemu_popcull():
    """The same with the SPRAY_POP_CULL traversal knob on (stale stack entries pruned by distance bytes)."""
    return _load_hostemu("libhostemu_popcull.so", ["-DSPRAY_POP_CULL=1"])


def make_rays(o, d, tmin=1e-3):
    o = np.asarray(o, np.float32).reshape(-1, 3)
    d = np.asarray(d, np.float32).reshape(-1, 3)
    r = np.zeros((o.shape[0], 8), np.float32)
    r[:, 0:3] = o
    r[:, 3] = tmin
    r[:, 4:7] = d
    r[:, 7] = np.inf
    return r


def random_rays(n, lo, hi, seed=0, spread=1.5):
    """Rays from a shell around [lo,hi] aimed at random points inside it."""
    rng = np.random.default_rng(seed)
    lo = np.asarray(lo, np.float32); hi = np.asarray(hi, np.float32)
    c = 0.5 * (lo + hi); e = 0.5 * (hi - lo)
    o = (c + rng.uniform(-1, 1, (n, 3)) * e * (1 + spread)).astype(np.float32)
    tgt = (c + rng.uniform(-1, 1, (n, 3)) * e).astype(np.float32)
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return make_rays(o, d.astype(np.float32))
