#!/usr/bin/env python3
"""CPU evaluation of BVH8 builder variants through the host emulation of core.cuh (tests/hostemu): nodes and
triangles visited per ray for camera rays and for incoherent rays.  usage: tools/bvh_quality.py "<extra g++ flags>" ..."""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from spray_b200 import scenes
from conftest import random_rays

def build(flags, tag):
    so = "/tmp/libhostemu_%s.so" % tag
    subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-mavx2", "-mfma",
                    "-Wno-unknown-pragmas"] + flags.split() + ["-o", so, os.path.join(ROOT, "tests/hostemu/hostemu.cpp")], check=True)
    E = C.CDLL(so)
    E.emu_build.restype = C.c_void_p
    E.emu_build.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    E.emu_num_nodes.restype = C.c_size_t; E.emu_num_nodes.argtypes = [C.c_void_p]
    E.emu_intersect.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return E

def run(E, verts, faces, rays):
    h = E.emu_build(verts.ctypes.data, len(verts), faces.ctypes.data, len(faces))
    n = len(rays); keys = np.zeros(n, np.uint64); nn, nt = C.c_uint64(), C.c_uint64()
    E.emu_intersect(h, 0, rays.ctypes.data, n, keys.ctypes.data, None, C.byref(nn), C.byref(nt))
    if os.environ.get("EMU_ORDER"):                  # measurement-only walk, children by entry distance
        E.emu_intersect_sorted(C.c_void_p(h), rays.ctypes.data, C.c_size_t(n), keys.ctypes.data, C.byref(nn), C.byref(nt))
    return E.emu_num_nodes(h), nn.value / n, nt.value / n, keys

wv, wf, _ = scenes.load_wavelet_mesh()
lo = np.zeros(3, np.float32); hi = np.full(3, 20, np.float32)
sv, sf, _ = scenes.soup_domain(0, 1 << 17, lo, hi)
sets = [("wavelet", np.ascontiguousarray(wv, np.float32), np.ascontiguousarray(wf, np.uint32), random_rays(40000, wv.min(0), wv.max(0), seed=1)),
        ("soup128k", np.ascontiguousarray(sv, np.float32), np.ascontiguousarray(sf, np.uint32), random_rays(40000, lo, hi, seed=2))]
ref = {}
for i, flags in enumerate(sys.argv[1:] or [""]):
    env = [t for t in flags.split() if t.startswith("EMU_")]          # e.g. EMU_BUILDER=sah (read by emu_build)
    for k in [k for k in os.environ if k.startswith("EMU_")]: del os.environ[k]
    for t in env: os.environ[t.split("=")[0]] = t.split("=", 1)[1]
    E = build(" ".join(t for t in flags.split() if not t.startswith("EMU_")), str(i))
    for name, v, f, rays in sets:
        nodes, nn, nt, keys = run(E, v, f, rays)
        same = "" if name not in ref else (" keys==variant0: %s" % np.array_equal(keys, ref[name]))
        ref.setdefault(name, keys)
        print("%-28s %-9s nodes %7d  nodes/ray %6.2f  tris/ray %6.2f%s" % (flags or "(default)", name, nodes, nn, nt, same))
