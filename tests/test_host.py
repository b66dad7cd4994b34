"""CPU tests of the host-side C++ mirror (host/): PLY and .spray loaders, caches, tiles, camera frame,
PPM writer and command-line parsing, through host/libsprayhost_test.so (no GPU needed).  Expected values
come from the golden fixtures decoded from the reference's own example files (tests/golden) and from the
oracle's restatement of the same reference functions."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from spray_b200 import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")


@pytest.fixture(scope="module")
def H():
    env = dict(os.environ)
    env.pop("CC", None); env.pop("CXX", None)
    subprocess.run(["make", "-C", HOST, "libsprayhost_test.so"], check=True, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    L = C.CDLL(os.path.join(HOST, "libsprayhost_test.so"))
    L.host_last_error.restype = C.c_char_p
    return L


def err(H):
    return H.host_last_error().decode()


def load_ply(H, path, nv_cap, nf_cap, with_colors=True):
    v = np.zeros(nv_cap * 3, np.float32); f = np.zeros(nf_cap * 3, np.uint32); c = np.zeros(nv_cap, np.uint32)
    nv = C.c_uint64(); nf = C.c_uint64()
    rc = H.host_ply_load(path.encode(), v.ctypes.data_as(C.c_void_p), C.c_uint64(v.size), f.ctypes.data_as(C.c_void_p),
                         C.c_uint64(f.size), c.ctypes.data_as(C.c_void_p) if with_colors else None, C.c_uint64(c.size),
                         C.byref(nv), C.byref(nf))
    return rc, v[:3 * nv.value].reshape(-1, 3), f[:3 * nf.value].reshape(-1, 3), c[:nv.value]


@pytest.mark.parametrize("binary", [True, False])
def test_ply_roundtrip_wavelet(H, tmp_path, binary):
    """The reference's examples/wavelet.ply content (golden npz) written in both dialects the reference reads
    (ply_loader.cc:206-324) comes back bit-exact, colours packed r<<16|g<<8|b."""
    verts, faces, rgb = scenes.load_wavelet_mesh()
    p = str(tmp_path / "w.ply")
    scenes.write_ply(p, verts, faces, rgb, binary=binary)
    nv = C.c_uint64(); nf = C.c_uint64(); hc = C.c_int()
    assert H.host_ply_header(p.encode(), C.byref(nv), C.byref(nf), C.byref(hc)) == 0, err(H)
    assert (nv.value, nf.value, hc.value) == (2840, 5480, 1)
    rc, v, f, c = load_ply(H, p, 2840, 5480)
    assert rc == 0, err(H)
    assert np.array_equal(v.view(np.uint32), verts.view(np.uint32))
    assert np.array_equal(f, faces)
    assert np.array_equal(c, rgb)


def test_ply_without_colors_and_short_index_types(H, tmp_path):
    verts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], np.float32)
    faces = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    p = str(tmp_path / "t.ply")
    scenes.write_ply(p, verts, faces, None, binary=True)
    rc, v, f, c = load_ply(H, p, 4, 2)
    assert rc == 0, err(H)
    assert np.array_equal(v, verts) and np.array_equal(f, faces.astype(np.uint32)) and (c == 0).all()
    # list uchar ushort: 2-byte indices are zero-extended (ply_loader.cc:296-303)
    hdr = ("ply\nformat binary_little_endian 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\n"
           "element face 2\nproperty list uchar ushort vertex_index\nend_header\n").encode()
    body = verts.tobytes() + b"".join(bytes([3]) + np.asarray(t, "<u2").tobytes() for t in faces)
    p2 = str(tmp_path / "t2.ply")
    open(p2, "wb").write(hdr + body)
    rc, v, f, c = load_ply(H, p2, 4, 2, with_colors=False)
    assert rc == 0, err(H)
    assert np.array_equal(f, faces.astype(np.uint32))


@pytest.mark.parametrize("mutation,needle", [
    (lambda s: s.replace(b"ply\n", b"plx\n", 1), "unknown file type"),
    (lambda s: s.replace(b"binary_little_endian", b"binary_big_endian"), "big endian"),
    (lambda s: s.replace(b"property float y\nproperty float z", b"property float z\nproperty float y"), "order"),
    (lambda s: s.replace(b"property float x", b"property double x"), "float"),
    (lambda s: s.replace(b"element face 2", b"element face 2\nelement edge 1"), "elements"),
    (lambda s: s[:-5], "truncated"),
])
def test_ply_malformed_is_fatal(H, tmp_path, mutation, needle):
    """Everything the reference CHECK-fails on (ply_loader.cc:114-176,330-440) is fatal here too."""
    verts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], np.float32)
    faces = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    p = str(tmp_path / "t.ply")
    scenes.write_ply(p, verts, faces, None, binary=True)
    data = mutation(open(p, "rb").read())
    open(p, "wb").write(data)
    rc, *_ = load_ply(H, p, 4, 2)
    assert rc == -1 and needle in err(H), err(H)


def test_ply_capacity_and_quads_are_fatal(H, tmp_path):
    verts, faces, rgb = scenes.load_wavelet_mesh()
    p = str(tmp_path / "w.ply")
    scenes.write_ply(p, verts, faces, rgb)
    rc, *_ = load_ply(H, p, 100, 5480)
    assert rc == -1 and "capacity" in err(H)
    hdr = ("ply\nformat ascii 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\n"
           "element face 1\nproperty list uchar int vertex_indices\nend_header\n")
    p2 = str(tmp_path / "q.ply")
    open(p2, "w").write(hdr + "0 0 0\n1 0 0\n1 1 0\n0 1 0\n4 0 1 2 3\n")
    rc, *_ = load_ply(H, p2, 4, 1)
    assert rc == -1 and "indices" in err(H)


def scene_load(H, path, ply_path=""):
    dom = np.zeros((512, 34), np.float64); names = C.create_string_buffer(512 * 256); li = np.zeros((8, 7), np.float64)
    nd = C.c_int(); nl = C.c_int()
    rc = H.host_scene_load(path.encode(), ply_path.encode(), 512, 8, dom.ctypes.data_as(C.c_void_p), names,
                           li.ctypes.data_as(C.c_void_p), C.byref(nd), C.byref(nl))
    fn = [names.raw[256 * i:256 * (i + 1)].split(b"\0")[0].decode() for i in range(max(nd.value, 0))]
    return rc, dom[:nd.value], fn, li[:nl.value]


def test_scene_loader_wavelets64(H, tmp_path):
    """A file with the tokens of examples/wavelets64/wavelets64.spray (golden JSON) parses to the same domain
    table the Python scene description and the oracle use: world bounds bit-exact (scene_loader.cc:352-357)."""
    sc = scenes.wavelets64()
    p = str(tmp_path / "w64.spray")
    scenes.write_spray_file(sc, p)
    rc, dom, fn, li = scene_load(H, p, "/data/ply")
    assert rc == 0, err(H)
    assert len(dom) == 64 and len(li) == 1
    gold = scenes.scene_from_golden_json()
    for i, d in enumerate(sc["domains"]):
        assert fn[i] == "/data/ply/wavelet.ply"
        assert np.array_equal(dom[i, 6:12].astype(np.float32), d["bounds"])
        m = np.eye(4, dtype=np.float32).T.reshape(16) if d["transform"] is None else d["transform"]
        assert np.array_equal(dom[i, 12:28].astype(np.float32), m)
        assert dom[i, 28] == 0 and tuple(dom[i, 29:32]) == (1.0, 1.0, 1.0)
        assert (dom[i, 32], dom[i, 33]) == (gold["domains"][i]["vertex"], gold["domains"][i]["face"])
        assert np.array_equal(dom[i, 0:6].astype(np.float32), np.asarray(gold["domains"][i]["bound"], np.float32))
    assert tuple(li[0]) == (0.0, 0.0, 500.0, 1000.0, 1.0, 1.0, 1.0)


def test_scene_loader_transform_order_and_materials(H, tmp_path):
    """scale / rotate / translate compose in file order with glm semantics; mtl and light variants."""
    p = str(tmp_path / "s.spray")
    open(p, "w").write(
        "# comment\nlight diffuse 0.5 0.6 0.7\nlight point 1 2 3 4 5 6\n\n"
        "domain\nfile a.ply\nmtl mirror 0.9 0.8 0.7\nbound -1 -2 -3 1 2 3\nface 10\nvertex 8\n"
        "translate 10 0 0\nscale 2 2 2\n\n"
        "domain\nfile b.ply\nmtl glass 1.0 1.5\nbound 0 0 0 1 1 1\nface 3\nvertex 5\nrotate z 90\ntranslate 1 0 0\n"
        "domain\nfile c.ply\nmtl transmission 1.0 1.3\nbound 0 0 0 1 1 1\nface 3\nvertex 5\n")
    rc, dom, fn, li = scene_load(H, p)
    assert rc == 0, err(H)
    assert fn == ["a.ply", "b.ply", "c.ply"]
    # translate then scale: M = T * S  ->  world = 2 * v + (10,0,0)
    assert np.allclose(dom[0, 6:12], [8, -4, -6, 12, 4, 6])
    assert dom[0, 28] == 1 and np.allclose(dom[0, 29:32], np.float32([0.9, 0.8, 0.7]))
    # rotate z 90 then translate (1,0,0): M = R * T -> point (x,y,z) -> R (x+1, y, z) = (-y, x+1, z)
    assert np.allclose(dom[1, 6:12], [0, 1, 0, -1, 2, 1], atol=1e-6)   # min/max corners only, NOT re-sorted
    assert dom[1, 28] == 2 and np.allclose(dom[1, 29:31], [1.0, 1.5])
    assert dom[2, 28] == 3
    assert li[0, 0] == 1 and np.allclose(li[0, 4:7], np.float32([0.5, 0.6, 0.7]))
    assert li[1, 0] == 0 and tuple(li[1, 1:7]) == (1, 2, 3, 4, 5, 6)


@pytest.mark.parametrize("text,needle", [
    ("file a.ply\ndomain\n", "before the first"),
    ("domain\nfoo 1\n", "unknown tag"),
    ("domain\nmtl plastic 1 1 1\n", "unknown material"),
    ("domain\nrotate w 10\n", "invalid axis"),
    ("domain\nbound 0 0 0 1 1\n", "bound"),
    ("light spot 0 0 0 1 1 1\ndomain\n", "unknown light"),
    ("# nothing\n", "no domains"),
])
def test_scene_loader_malformed_is_fatal(H, tmp_path, text, needle):
    p = str(tmp_path / "bad.spray")
    open(p, "w").write(text)
    rc, *_ = scene_load(H, p)
    assert rc == -1 and needle in err(H), err(H)


def test_lru_and_infinite_cache_sequences(H, oracle):
    """Same hit/miss/block sequences as the oracle's restatement of lru_cache.cc:65-171."""
    rng = np.random.default_rng(3)
    for nd, cap in [(8, 3), (64, 16), (5, -1), (5, 9), (16, 1)]:
        req = rng.integers(0, nd, 400).astype(np.int32)
        out = np.zeros(len(req), np.int32); capacity = C.c_int()
        assert H.host_cache_replay(1, nd, cap, req.ctypes.data_as(C.c_void_p), len(req), out.ctypes.data_as(C.c_void_p),
                                   C.byref(capacity)) == 0, err(H)
        hits = np.zeros(len(req), np.int32); blocks = np.zeros(len(req), np.int32)
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))
        ecap = nd if cap < 0 or cap >= nd else cap          # LruCache::init, lru_cache.cc:50-54
        oracle.lib().orc_lru_sequence(nd, ecap, ip(req), len(req), ip(hits), ip(blocks))
        assert np.array_equal(out & 0xFFFF, blocks) and np.array_equal(out >> 30, hits)
        assert capacity.value == ecap
    req = np.array([3, 1, 3, 0, 1], np.int32); out = np.zeros(5, np.int32); capacity = C.c_int()
    assert H.host_cache_replay(0, 4, -1, req.ctypes.data_as(C.c_void_p), 5, out.ctypes.data_as(C.c_void_p), C.byref(capacity)) == 0
    assert list(out & 0xFFFF) == [3, 1, 3, 0, 1] and list(out >> 30) == [0, 0, 1, 0, 1]
    assert H.host_cache_replay(0, 4, 2, req.ctypes.data_as(C.c_void_p), 5, out.ctypes.data_as(C.c_void_p), C.byref(capacity)) == -1
    assert H.host_cache_replay(1, 4, 0, req.ctypes.data_as(C.c_void_p), 5, out.ctypes.data_as(C.c_void_p), C.byref(capacity)) == -1


@pytest.mark.parametrize("w,h,spp,nr,cap", [(2048, 2048, 1, 1, 1 << 20), (2048, 2048, 1, 4, 1 << 20), (512, 512, 4, 2, 1 << 20),
                                            (4096, 4096, 4, 8, 1 << 20), (100, 37, 3, 3, 500), (7, 5, 1, 16, 1 << 20)])
def test_tiles_match_reference_restatement(H, w, h, spp, nr, cap):
    for rank in range(nr):
        xy = np.zeros((4096, 4), np.int32); nt = C.c_int(); big = np.zeros(4, np.int32)
        assert H.host_tiles(w, h, spp, nr, rank, C.c_int64(cap), 4096, xy.ctypes.data_as(C.c_void_p), C.byref(nt),
                            big.ctypes.data_as(C.c_void_p)) == 0, err(H)
        got = [tuple(int(v) for v in r) for r in xy[:nt.value]]
        assert got == scenes.image_schedule_tiles(w, h, spp, nr, rank, cap)
        if got:
            assert tuple(big) == max(got, key=lambda t: (t[2] * t[3], -got.index(t)))


def test_stripes(H):
    out = np.zeros(4, np.int32)
    tile = np.array([0, 10, 100, 35], np.int32)
    rows = []
    for r in range(4):
        H.host_stripes(1, 4, r, tile.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        rows.append(tuple(out))
    assert rows == [(0, 10, 100, 8), (0, 18, 100, 8), (0, 26, 100, 8), (0, 34, 100, 11)]   # tile.cc:204-226
    H.host_stripes(0, 200, 150, tile.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    assert (out[2], out[3]) == (0, 0)                                                      # rank beyond the width


def test_camera_frame_bit_exact(H, oracle):
    from spray_b200 import capi
    p = scenes.WAVELETS64_CAMERA
    for (w, h, fov) in [(2048, 2048, 90.0), (1024, 768, 60.0), (333, 517, 37.5)]:
        cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], fov, w, h)
        out = capi.Camera()
        f3 = lambda v: np.asarray(v, np.float32).ctypes.data_as(C.c_void_p)
        assert H.host_camera(f3(p["pos"]), f3(p["lookat"]), f3(p["up"]), C.c_float(fov), w, h, C.byref(out)) == 0, err(H)
        for k in ("pos", "lowerleft", "wvec", "hvec"):
            assert np.array_equal(np.asarray(getattr(out, k)[:], np.float32).view(np.uint32),
                                  np.asarray(cd[k], np.float32).view(np.uint32)), k
        assert (out.image_w, out.image_h) == (w, h)


def test_ppm_writer(H, tmp_path):
    """image.h:183-204: P3, maxval 1023, rows bottom-up, truncation after clamping."""
    img = np.zeros((2, 3, 4), np.float32)
    img[0, 0, :3] = [0.0, 0.5, 1.0]; img[0, 1, :3] = [2.0, -1.0, 0.9999]; img[1, 2, :3] = [0.25, 0.001, 1e-9]
    p = str(tmp_path / "o.ppm")
    assert H.host_write_ppm(p.encode(), 3, 2, img.ctypes.data_as(C.c_void_p)) == 0, err(H)
    toks = open(p).read().split()
    assert toks[:4] == ["P3", "3", "2", "1023"]
    vals = np.array(toks[4:], np.int64).reshape(2, 3, 3)
    q = lambda v: int(np.float32(min(max(np.float32(v) * np.float32(1023.0), np.float32(0)), np.float32(1023))))
    assert vals[0, 2].tolist() == [q(0.25), q(0.001), 0]           # image row 1 is written first
    assert vals[1, 0].tolist() == [0, q(0.5), 1023] and vals[1, 1].tolist() == [1023, 0, q(0.9999)]


def test_config_parse_reference_command_line(H):
    """The flags of examples/wavelets64/wavelets64.sh:141-163 (film mode)."""
    argv = ["spray_ooc", "--nthreads", "2", "-w", "512", "-h", "512", "--frames", "10", "--mode", "film", "--cache-size", "-1",
            "--partition", "image", "--camera", "90.172180", "84.141418", "82.480225", "30", "28.649426", "30",
            "--pixel-samples", "1", "--bounces", "1", "--blinn", "0.4", "0.4", "0.4", "10", "--ao-samples", "2",
            "--ply-path", "/x/y", "-o", "out.ppm", "scene.spray"]
    arr = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
    out = np.zeros(40, np.float64); model = C.create_string_buffer(256); ply = C.create_string_buffer(256); o = C.create_string_buffer(256)
    assert H.host_config_parse(len(argv), arr, out.ctypes.data_as(C.c_void_p), model, ply, o) == 0, err(H)
    f = np.float32
    assert out[:3].tolist() == [512, 512, 1]
    assert out[3:9].astype(np.float32).tolist() == [f(90.172180), f(84.141418), f(82.480225), f(30), f(28.649426), f(30)]
    assert out[9:12].tolist() == [0, 1, 0]
    assert out[12] == 90 and out[13] == 10 and out[15] == 1 and out[16] == 0      # fov, frames, bounces, partition image
    assert out[18] == 1 and out[19] == -1 and out[20] == 2 and out[21] == 0       # film, cache -1, ao samples, ao mode
    assert out[22] == 1 and out[23] == 1 << 20 and out[24] == 2                   # spp, max samples, nthreads
    assert out[26] == 10 and out[27:30].astype(np.float32).tolist() == [f(0.4)] * 3
    assert (model.value, ply.value, o.value) == (b"scene.spray", b"/x/y", b"out.ppm")
    bad = ["spray_ooc", "--mode", "film", "scene.txt"]
    arr = (C.c_char_p * len(bad))(*[a.encode() for a in bad])
    assert H.host_config_parse(len(bad), arr, out.ctypes.data_as(C.c_void_p), model, ply, o) == -1
