"""GPU tests of spraygpu_scene_commit's device-side collapse (build.cu, round 2): the cooperative 8-lanes-per-node kernel
must build THE SAME compressed 8-wide tree as collapse_select / encode_node8 run serially on the CPU (tests/hostemu,
the code the first round's kernel executed with one thread per node): node for node, byte for byte, up to the order in
which sibling subtrees were allocated.  Plus the enqueue-only commit (spraygpu_scene_commit_async)."""
import numpy as np
import pytest

from spray_b200 import capi, scenes
from test_hostemu import adversarial_meshes
from conftest import random_rays

pytestmark = pytest.mark.gpu


def _emu_bvh(E, verts, faces):
    verts = np.ascontiguousarray(verts, np.float32); faces = np.ascontiguousarray(faces, np.uint32)
    h = E.emu_build(verts.ctypes.data, len(verts), faces.ctypes.data, len(faces))
    assert h
    nn = E.emu_num_nodes(h)
    nodes = np.zeros((nn, 20), np.uint32); tris = np.zeros((len(faces), 12), np.float32)
    E.emu_get_bvh(h, nodes.ctypes.data, tris.ctypes.data)
    depth = E.emu_depth(h)
    E.emu_destroy(h)
    return nodes, tris, depth


def _same_tree(na, ta, nb, tb):
    """Walk both trees from the root in slot order; everything but child_base / tri_base (allocation order) must be equal."""
    ta = ta.view(np.uint32); tb = tb.view(np.uint32)
    stack = [(0, 0)]
    visited = 0
    while stack:
        a, b = stack.pop()
        qa, qb = na[a], nb[b]
        visited += 1
        if not (np.array_equal(qa[0:4], qb[0:4]) and np.array_equal(qa[6:20], qb[6:20])):
            return "node words differ at (%d, %d): %s vs %s" % (a, b, qa, qb)
        imask = int(qa[3]) >> 24
        meta = [(int(qa[6 + (s >> 2)]) >> (8 * (s & 3))) & 0xff for s in range(8)]
        rank = 0
        for s in range(8):
            m = meta[s]
            if m == 0:
                continue
            if (imask >> s) & 1:
                stack.append((int(qa[4]) + rank, int(qb[4]) + rank))
                rank += 1
            else:
                cnt = {1: 1, 3: 2, 7: 3}[m >> 5]
                off = m & 31
                for k in range(cnt):
                    ia, ib = int(qa[5]) + off + k, int(qb[5]) + off + k
                    if not np.array_equal(ta[ia], tb[ib]):
                        return "triangle record differs under node (%d, %d) slot %d" % (a, b, s)
    return None if visited == len(na) == len(nb) else "visited %d of %d / %d nodes" % (visited, len(na), len(nb))


def _meshes():
    sc = scenes.wavelets64(1)
    d = sc["domains"][0]
    out = [("wavelet", d["verts"], d["faces"])]
    for nt in (2, 3, 4, 9, 1000, 1 << 14):
        v, f, _ = scenes.soup_domain(11 + nt, nt, (0, 0, 0), (20, 20, 20))
        out.append(("soup%d" % nt, v, f))
    for i, (v, f) in enumerate(adversarial_meshes()):
        out.append(("adversarial%d" % i, v, f))
    return out


def test_device_collapse_builds_the_emulated_tree(hostemu):
    dev = capi.Device(0)
    for name, v, f in _meshes():
        lo, hi = np.asarray(v).min(0) - 1, np.asarray(v).max(0) + 1
        dev.set_domains([np.concatenate([lo, hi]).astype(np.float32)], [(0, (1, 1, 1))])
        s = dev.new_scene()
        s.set_mesh(v, f, None)
        s.commit()
        inf = s.info()
        ng, tg = s.bvh()
        ne, te, depth = _emu_bvh(hostemu, v, f)
        assert inf.num_nodes == len(ne), (name, inf.num_nodes, len(ne))
        assert inf.bvh_depth == depth, (name, inf.bvh_depth, depth)
        # every triangle exactly once in the leaf-order records
        prims = np.sort(tg.view(np.uint32)[:, 3])
        assert np.array_equal(prims, np.arange(len(f), dtype=np.uint32)), name
        err = _same_tree(ng, tg, ne, te)
        assert err is None, (name, err)
        s.close()
    dev.close()


def test_commit_async_then_trace(oracle):
    """spraygpu_scene_commit_async: several builds enqueued back to back without a host wait, bound at once; the first
    trace call waits for them (upload_domain_table) and the hits are the oracle's."""
    sc = scenes.wavelets64(4)
    dev = capi.Device(0)
    doms = sc["domains"]
    dev.set_domains([d["bounds"] for d in doms], [d.get("mtl", (0, (1, 1, 1))) for d in doms])
    held = []
    for i, d in enumerate(doms):
        s = dev.new_scene()
        s.set_mesh(d["verts"], d["faces"], d.get("rgb"))
        s.set_transform(d.get("transform"))
        s.commit(wait=False)
        dev.bind(i, s)
        held.append(s)
    lo, hi = scenes.scene_bounds(sc)[:3], scenes.scene_bounds(sc)[3:]
    rays = random_rays(20000, lo, hi, seed=9)
    keys = np.zeros(len(rays), np.uint64)
    dev.intersect_domains(rays, len(rays), keys, None)
    ko, _ = oracle.Scene(doms).intersect(rays)
    assert np.array_equal(keys, ko)
    assert all(s.info().num_nodes > 0 and s.info().build_ms > 0 for s in held)
    # re-commit into the same blocks (an out-of-core cache block that is re-loaded), again without waiting
    for s in held:
        s.commit(wait=False)
    dev.intersect_domains(rays, len(rays), keys, None)
    assert np.array_equal(keys, ko)
    dev.close()


def test_normals_of_high_valence_vertices_bit_exact(oracle):
    """computeNormals (trimesh_buffer.cc:267-326) sums a vertex's face normals in increasing face order.  The commit
    gathers each vertex's faces with atomics (arbitrary order) and sorts them per vertex: in local memory up to 32
    faces, with a block-wide bitonic network beyond — a fan whose hub has 5 000 faces, rims of 32 / 33 / 100 faces, a
    face that names the same vertex twice."""
    rng = np.random.default_rng(12)
    verts = [np.zeros(3, np.float32)]
    faces = []

    def fan(hub, nfaces, centre, radius):
        base = len(verts)
        ang = np.sort(rng.uniform(0, 2 * np.pi, nfaces + 1))
        for a in ang:
            verts.append(np.asarray(centre, np.float32) + np.float32(radius) * np.array([np.cos(a), np.sin(a), rng.uniform(-0.2, 0.2)], np.float32))
        for k in range(nfaces):
            faces.append((hub, base + k, base + k + 1))

    fan(0, 5000, (0, 0, 0), 7.0)
    for nf_, c in ((32, (20, 0, 0)), (33, (0, 20, 0)), (100, (20, 20, 0)), (1, (40, 0, 0))):
        verts.append(np.asarray(c, np.float32) + np.float32(0.5))
        fan(len(verts) - 1, nf_, c, 3.0)
    faces.append((0, 0, 5))                                  # degenerate: the hub twice
    v = np.asarray(verts, np.float32)
    f = np.asarray(faces, np.uint32)
    perm = rng.permutation(len(f))                           # face order = summation order: shuffle it
    f = np.ascontiguousarray(f[perm])
    lo, hi = v.min(0) - 1, v.max(0) + 1
    dom = dict(verts=v, faces=f, rgb=None, transform=None, bounds=np.concatenate([lo, hi]).astype(np.float32))
    osc = oracle.Scene([dom])
    dev = capi.Device(0)
    dev.set_domains([dom["bounds"]], [(0, (1, 1, 1))])
    s = dev.new_scene()
    s.set_mesh(v, f, None)
    for _ in range(2):                                       # twice: the atomics order the lists differently each time
        s.commit()
        assert np.array_equal(s.normals(), osc.normals(0))
    dev.bind(0, s)
    rays = random_rays(5000, lo, hi, seed=4)
    keys = np.zeros(len(rays), np.uint64)
    dev.intersect_domains(rays, len(rays), keys, None)
    assert np.array_equal(keys, osc.intersect(rays)[0])
    dev.close()
