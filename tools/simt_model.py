#!/usr/bin/env python3
"""CPU model of the warp-synchronous traversal loop (tests/hostemu emu_simt): lanes per executed step for the
product's 32-ray batches vs an ideal refill stream, on coherent (camera) and incoherent rays.
usage: tools/simt_model.py [-DSPRAY_POP_CULL=1]"""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from spray_b200 import scenes
from conftest import random_rays, make_rays

so = "/tmp/libhostemu_simt.so"
subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-mavx2", "-mfma",
                "-Wno-unknown-pragmas"] + sys.argv[1:] + ["-o", so, os.path.join(ROOT, "tests/hostemu/hostemu.cpp")], check=True)
E = C.CDLL(so)
E.emu_build.restype = C.c_void_p
E.emu_build.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
E.emu_simt.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
CN, CT = (281.0 if "-DSPRAY_POP_CULL=1" in sys.argv else 257.0), 35.0     # SASS instructions of one node step / one triangle step (r02c build)
CR = 150.0               # assumed cost of one refill execution (park the hits, pop + set up the next rays)

def camera_rays(lo, hi, w=256, h=256):
    c = 0.5 * (lo + hi); e = (hi - lo)
    eye = c + np.array([0, 0, 1.6], np.float32) * e
    x, y = np.meshgrid(np.linspace(-0.6, 0.6, w), np.linspace(-0.6, 0.6, h))
    # 8x4 pixel blocks per warp, as the kernel hands them out
    idx = np.arange(w * h).reshape(h // 4, 4, w // 8, 8).transpose(0, 2, 1, 3).reshape(-1)
    tgt = np.stack([c[0] + x.ravel() * e[0], c[1] + y.ravel() * e[1], np.full(w * h, c[2])], 1)[idx]
    return make_rays(np.tile(eye, (w * h, 1)), tgt - eye)

wv, wf, _ = scenes.load_wavelet_mesh()
wv = np.ascontiguousarray(wv, np.float32); wf = np.ascontiguousarray(wf, np.uint32)
lo = np.zeros(3, np.float32); hi = np.full(3, 20, np.float32)
sv, sf, _ = scenes.soup_domain(0, 1 << 17, lo, hi)
sv = np.ascontiguousarray(sv, np.float32); sf = np.ascontiguousarray(sf, np.uint32)
cases = [("wavelet camera", wv, wf, camera_rays(wv.min(0), wv.max(0))),
         ("wavelet incoherent", wv, wf, random_rays(65536, wv.min(0), wv.max(0), seed=1)),
         ("soup128k camera", sv, sf, camera_rays(lo, hi)),
         ("soup128k incoherent", sv, sf, random_rays(65536, lo, hi, seed=2))]
print("%-22s %-10s %9s %9s %9s %9s %9s %12s" % ("rays", "policy", "node ex", "lanes/ex", "tri ex", "lanes/ex", "refills", "warp-instr/ray"))
for name, v, f, rays in cases:
    h = E.emu_build(v.ctypes.data, len(v), f.ctypes.data, len(f))
    base = None
    for refill in (0, 1, 4, 8, 16, 24, -4, -8, -12, -16):
        out = np.zeros(7, np.uint64)
        E.emu_simt(C.c_void_p(h), rays.ctypes.data, C.c_size_t(len(rays)), refill, out.ctypes.data)
        ne, nl, te, tl, npass, rex, rl = [float(x) for x in out]
        cost = (ne * CN + te * CT + rex * CR) / len(rays)
        base = base or cost
        print("%-22s %-10s %9.0f %9.1f %9.0f %9.1f %9.0f %12.1f  (%.2fx)  passes/32 rays %6.1f" % (name, ("refill>=%d" % refill if refill > 0 else "cut<%d" % -refill) if refill else "batch32", ne, nl / max(ne, 1), te, tl / max(te, 1), rex, cost, base / cost, npass * 32 / len(rays)))
