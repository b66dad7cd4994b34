"""spraygpu_compute_normals_static — the per-vertex normals stage for callers that may not allocate.

Who needs it: SpRay's in-situ mode puts the renderer inside a running simulation.  That host has normally given the
whole of HBM to its own pool, so a cudaMalloc from the render plug-in fails or takes memory the solver budgeted for,
and cudaMalloc / cudaFree synchronise the device and stall the solver's streams.  The entry point therefore keeps its
work space (valences, per-vertex face lists) in a fixed array inside the library image and creates nothing.

What is checked, through the C ABI only: the normals are bit-identical (integer/byte bar of the other build tests) to
the regular path's (spraygpu_scene_set_mesh + spraygpu_scene_commit + spraygpu_scene_get_normals) on calls made
back to back over the SAME shared array with different meshes — the second call must not see the first one's counters or
lists — that free device memory does not move across a call, and that oversize / host / malformed inputs are refused
with a readable error.  The file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must
not hide the parity tests.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from spray_b200 import capi

pytestmark = pytest.mark.gpu


def _fan_mesh(seed, hub_faces):
    """A hub with `hub_faces` incident faces (heavy vertex: block-wide sort), rims of 32 / 33 / 100 faces (the boundary
    between the per-thread and the per-CTA path), a degenerate face, faces shuffled (the face order IS the summation order)."""
    rng = np.random.default_rng(seed)
    verts = [np.zeros(3, np.float32)]
    faces = []

    def fan(hub, nfaces, centre, radius):
        base = len(verts)
        ang = np.sort(rng.uniform(0, 2 * np.pi, nfaces + 1))
        for a in ang:
            verts.append(np.asarray(centre, np.float32) + np.float32(radius) * np.array([np.cos(a), np.sin(a), rng.uniform(-0.2, 0.2)], np.float32))
        for k in range(nfaces):
            faces.append((hub, base + k, base + k + 1))

    fan(0, hub_faces, (0, 0, 0), 7.0)
    for nf_, c in ((32, (20, 0, 0)), (33, (0, 20, 0)), (100, (20, 20, 0)), (1, (40, 0, 0))):
        verts.append(np.asarray(c, np.float32) + np.float32(0.5))
        fan(len(verts) - 1, nf_, c, 3.0)
    faces.append((0, 0, 5))
    v = np.asarray(verts, np.float32)
    f = np.asarray(faces, np.uint32)
    return v, np.ascontiguousarray(f[rng.permutation(len(f))])


def _grid_mesh(n, seed):
    """n x n height field, two triangles per cell (valence <= 6 everywhere: the per-thread path only)."""
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:n, 0:n]
    v = np.stack([xs, ys, rng.uniform(-1, 1, (n, n))], -1).reshape(-1, 3).astype(np.float32)
    i = (ys[:-1, :-1] * n + xs[:-1, :-1]).reshape(-1)
    f = np.concatenate([np.stack([i, i + 1, i + n], 1), np.stack([i + 1, i + n + 1, i + n], 1)]).astype(np.uint32)
    return v, np.ascontiguousarray(f[rng.permutation(len(f))])


def _regular_path(dev, v, f):
    s = dev.new_scene()
    s.set_mesh(v, f, None)
    s.commit()
    out = s.normals()
    s.close()
    return out


def _static_path(torch, v, f, stream=None):
    dv = torch.from_numpy(v).cuda()
    df = torch.from_numpy(f.view(np.int32)).cuda()
    dn = torch.full((v.shape[0], 3), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    capi.compute_normals_static(dv, df, dn, stream)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return dn.cpu().numpy(), free0, free1


def test_static_normals_match_the_regular_path_on_consecutive_calls():
    import torch
    torch.cuda.set_device(0)
    dev = capi.Device(0)
    big = _fan_mesh(12, 5000)        # 5 166 faces, a 5 001-face hub
    small = _fan_mesh(5, 40)         # fewer vertices and faces than the call before it: stale cursors would show
    grid = _grid_mesh(181, 3)        # 32 761 vertices / 64 800 faces
    want = [_regular_path(dev, *m) for m in (big, small, grid)]
    dev.close()

    # first call on this GPU (also the one that resolves the scratch array's address), then three more, back to back,
    # all over the same array; the last one on a stream of the caller's
    side = torch.cuda.Stream()
    order = [(big, want[0], None), (small, want[1], None), (grid, want[2], None), (big, want[0], side.cuda_stream)]
    for k, (mesh, ref, stream) in enumerate(order):
        got, free0, free1 = _static_path(torch, *mesh, stream=stream)
        assert not np.isnan(got).any()
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), "call %d differs from spraygpu_scene_commit's normals" % k
        if k > 0:                    # call 0 may load the kernels' module lazily; afterwards nothing may be taken
            assert free1 >= free0, "device memory shrank by %d B across a call that must not allocate" % (free0 - free1)


def test_static_normals_refuse_what_they_cannot_hold():
    import torch
    torch.cuda.set_device(0)
    L = capi.lib()
    v, f = _fan_mesh(1, 10)
    dv = torch.from_numpy(v).cuda()
    df = torch.from_numpy(f.view(np.int32)).cuda()
    dn = torch.zeros((v.shape[0], 3), dtype=torch.float32, device="cuda")
    nv, nf = v.shape[0], f.shape[0]
    call = L.spraygpu_compute_normals_static
    # beyond the built-in array: a clear limit error naming the maximum, nothing launched
    assert call(dv.data_ptr(), capi.STATIC_MAX_VERTS + 1, df.data_ptr(), nf, dn.data_ptr(), None) == -4
    assert b"exceed" in L.spraygpu_last_error() and str(capi.STATIC_MAX_VERTS).encode() in L.spraygpu_last_error()
    assert call(dv.data_ptr(), nv, df.data_ptr(), capi.STATIC_MAX_FACES + 1, dn.data_ptr(), None) == -4
    # host arrays would need staging memory
    assert call(v.ctypes.data, nv, df.data_ptr(), nf, dn.data_ptr(), None) == -2
    assert b"device memory" in L.spraygpu_last_error()
    # an index past the last vertex must not become a write outside the array
    bad = f.copy()
    bad[3, 1] = nv
    dbad = torch.from_numpy(bad.view(np.int32)).cuda()
    assert call(dv.data_ptr(), nv, dbad.data_ptr(), nf, dn.data_ptr(), None) == -2
    assert b"out of range" in L.spraygpu_last_error()
    assert call(None, nv, df.data_ptr(), nf, dn.data_ptr(), None) == -2
    # and the array is still good for the next caller
    dev = capi.Device(0)
    ref = _regular_path(dev, v, f)
    dev.close()
    got, _, _ = _static_path(torch, v, f)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
