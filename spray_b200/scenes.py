"""Scene descriptions for tests and benchmarks (host-side, numpy only).

A scene is a list of domain dicts {verts f32[nv,3] (object space), faces u32[nf,3], rgb u32[nv] packed
0x00RRGGBB, transform f32[16] column-major or None, bounds f32[6] world AABB, mtl (type, params3)} plus
lights [(type, position3, radiance3)] — the content of a `.spray` file (reference
src/io/scene_loader.cc:42-358) after PLY loading (src/io/ply_loader.cc).
"""
import json
import math
import os

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(_ROOT, "tests", "golden")

MTL_DIFFUSE, MTL_MIRROR, MTL_GLASS, MTL_TRANSMISSION = 0, 1, 2, 3
LIGHT_POINT, LIGHT_DIFFUSE = 0, 1

f32 = np.float32


def pack_rgb(rgb_u8):
    """r<<16 | g<<8 | b  (reference src/io/ply_loader.cc:247)."""
    c = np.asarray(rgb_u8, dtype=np.uint32)
    return (c[:, 0] << 16) | (c[:, 1] << 8) | c[:, 2]


def load_wavelet_mesh(path=None):
    """The reference's examples/wavelet.ply decoded into arrays (tests/golden/make_fixtures.py)."""
    z = np.load(path or os.path.join(GOLDEN, "wavelet_mesh.npz"))
    return z["verts"].astype(np.float32), z["faces"].astype(np.uint32), pack_rgb(z["rgb"])


def translate_mat4(t):
    """glm::translate(mat4(1), t), column-major (reference src/io/scene_loader.cc:178-186)."""
    m = np.eye(4, dtype=np.float32)
    m[:3, 3] = np.asarray(t, np.float32)
    return np.ascontiguousarray(m.T).reshape(16)  # column-major: element (row r, col c) at c*4+r


def is_identity(m):
    return m is None or np.array_equal(np.asarray(m, np.float32).reshape(16), np.eye(4, dtype=np.float32).reshape(16))


def world_bounds(obj_bounds, m):
    """world AABB = transform * object min/max corners only (scene_loader.cc:352-357; wrong under rotation,
    reproduced on purpose).  glm mat4*vec4 order: (m0*x + m1*y) + (m2*z + m3*1)."""
    b = np.asarray(obj_bounds, np.float32)
    if m is None:
        return b.copy()
    m = np.asarray(m, np.float32).reshape(16)
    out = np.zeros(6, np.float32)
    for k in range(2):
        x, y, z = b[3 * k], b[3 * k + 1], b[3 * k + 2]
        for r in range(3):
            out[3 * k + r] = f32(f32(f32(m[r] * x) + f32(m[4 + r] * y)) + f32(f32(m[8 + r] * z) + f32(m[12 + r] * f32(1))))
    return out


def wavelet_single(light=True):
    """Config C1: examples/wavelet.ply as one domain + `light point 0 500 1000 1 1 1`."""
    verts, faces, rgb = load_wavelet_mesh()
    bounds = np.array([-10, -10, -10, 10, 9.324713, 10], np.float32)
    dom = dict(verts=verts, faces=faces, rgb=rgb, transform=None, bounds=bounds, mtl=(MTL_DIFFUSE, (1.0, 1.0, 1.0)))
    lights = [(LIGHT_POINT, (0.0, 500.0, 1000.0), (1.0, 1.0, 1.0))] if light else []
    return dict(domains=[dom], lights=lights)


def wavelets64(ndomains=64):
    """Config C2: 4x4x4 translated copies of wavelet.ply — the layout of
    examples/wavelets64/wavelets64.spray (checked against tests/golden/wavelets64_scene.json)."""
    verts, faces, rgb = load_wavelet_mesh()
    ob = np.array([-10, -10, -10, 10, 9.324713, 10], np.float32)
    doms = []
    for i in range(ndomains):
        ix, iy, iz = i % 4, (i // 4) % 4, i // 16
        # the literal values printed in the reference's scene file (atof'ed there, scene_loader.cc:178-186)
        t = (f32((0.0, 20.0, 40.0, 60.0)[ix]), f32((0.0, 19.324713, 38.649426, 57.974139)[iy]),
             f32((0.0, 20.0, 40.0, 60.0)[iz]))
        m = None if i == 0 else translate_mat4(t)
        doms.append(dict(verts=verts, faces=faces, rgb=rgb, transform=m, bounds=world_bounds(ob, m),
                         mtl=(MTL_DIFFUSE, (1.0, 1.0, 1.0)), file="wavelet.ply", translate=[float(x) for x in t]))
    return dict(domains=doms, lights=[(LIGHT_POINT, (0.0, 500.0, 1000.0), (1.0, 1.0, 1.0))])


WAVELETS64_CAMERA = dict(pos=(90.172180, 84.141418, 82.480225), lookat=(30.0, 28.649426, 30.0), up=(0.0, 1.0, 0.0),
                         vfov=90.0)  # examples/wavelets64/wavelets64.sh:141
WAVELETS64_BLINN = dict(ks=(0.4, 0.4, 0.4), shininess=10.0)  # wavelets64.sh:146


def scene_from_golden_json(path=None):
    """The reference's own wavelets64.spray as parsed by tests/golden/make_fixtures.py."""
    with open(path or os.path.join(GOLDEN, "wavelets64_scene.json")) as f:
        return json.load(f)


def soup_domain(seed, ntris, cell_lo, cell_hi):
    """Synthetic triangle soup for one domain (config C3 family): centres uniform in the cell, two edge
    vectors of length U(0.5,2)*h with h = extent / ntris^(1/3), vertices clipped to the cell so the domain
    AABB is the cell; vertex colours hashed from the vertex index.  numpy Philox keyed by `seed`."""
    rng = np.random.Generator(np.random.Philox(key=0x5EED0000 + int(seed)))
    lo = np.asarray(cell_lo, np.float32); hi = np.asarray(cell_hi, np.float32)
    ext = hi - lo
    h = float(ext.max()) / float(ntris) ** (1.0 / 3.0)
    c = lo + rng.random((ntris, 3), dtype=np.float32) * ext

    def edge():
        d = rng.standard_normal((ntris, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32) + f32(1e-20)
        return d * (f32(0.5) + f32(1.5) * rng.random((ntris, 1), dtype=np.float32)) * f32(h)

    e1, e2 = edge(), edge()
    v = np.stack([c, c + e1, c + e2], axis=1).reshape(-1, 3)
    v = np.minimum(np.maximum(v, lo), hi).astype(np.float32)
    faces = np.arange(3 * ntris, dtype=np.uint32).reshape(ntris, 3)
    idx = np.arange(3 * ntris, dtype=np.uint32)
    hsh = (idx * np.uint32(2654435761)) ^ (idx >> np.uint32(13))
    rgb = (np.uint32(64) + (hsh & np.uint32(0x7F))) << np.uint32(16) | (np.uint32(64) + ((hsh >> np.uint32(8)) & np.uint32(0x7F))) << np.uint32(8) | (np.uint32(64) + ((hsh >> np.uint32(16)) & np.uint32(0x7F)))
    return v, faces, rgb.astype(np.uint32)


def soup_scene(grid=(4, 4, 4), tris_per_domain=1 << 14, cell=20.0, mtl=(MTL_DIFFUSE, (1.0, 1.0, 1.0)), seed0=0):
    """Config C3-style scene: grid of cells, one soup domain each (geometry already in world space)."""
    doms = []
    gx, gy, gz = grid
    for i in range(gx * gy * gz):
        ix, iy, iz = i % gx, (i // gx) % gy, i // (gx * gy)
        lo = np.array([ix * cell, iy * cell, iz * cell], np.float32)
        hi = lo + f32(cell)
        v, f, rgb = soup_domain(seed0 + i, tris_per_domain, lo, hi)
        doms.append(dict(verts=v, faces=f, rgb=rgb, transform=None, bounds=np.concatenate([lo, hi]).astype(np.float32), mtl=mtl))
    return dict(domains=doms, lights=[(LIGHT_POINT, (0.0, 500.0, 1000.0), (1.0, 1.0, 1.0))])


def soup_h(ntris, cell_lo, cell_hi):
    """Mean triangle spacing of a soup cell: extent / ntris^(1/3), as the float both generators are handed."""
    ext = np.asarray(cell_hi, np.float32) - np.asarray(cell_lo, np.float32)
    return np.float32(float(ext.max()) / float(ntris) ** (1.0 / 3.0))


def soup_domain_device(seed, ntris, cell_lo, cell_hi, device="cuda"):
    """The soup domain of configs C3 / C4 generated ON THE GPU by the library's counter-based generator
    (spraygpu_gen_soup: Philox-4x32-10, key 0x5EED0000 + seed, counter = primID — SURVEY 8d): 64 x 2^20 ... 64 x 2^23
    triangles per GPU appear in milliseconds and never touch the host, and the CPU checker regenerates the SAME triangles
    from the same function (bit-identical, tests/test_gpu_sizes.py), so the benched scene can be handed to it.
    Returns CUDA tensors (verts f32[3n,3], faces i32[n,3] holding u32 values, rgb i32[3n])."""
    import torch
    from . import capi
    v = torch.empty((3 * ntris, 3), dtype=torch.float32, device=device)
    faces = torch.empty((ntris, 3), dtype=torch.int32, device=device)
    rgb = torch.empty(3 * ntris, dtype=torch.int32, device=device)
    with torch.cuda.device(v.device):
        capi.gen_soup(0x5EED0000 + int(seed), ntris, cell_lo, cell_hi, soup_h(ntris, cell_lo, cell_hi), v, faces, rgb,
                      torch.cuda.current_stream().cuda_stream)
    return v, faces, rgb


def soup_scene_device(grid=(4, 4, 4), tris_per_domain=1 << 20, cell=20.0, mtl=(MTL_DIFFUSE, (1.0, 1.0, 1.0)), seed0=0,
                      only=None, lazy=False):
    """soup_scene with device-resident meshes: domains carry `verts/faces/rgb` as CUDA tensors (or None for the
    domains not in `only` — another rank's share); capi.Device.load_scene uploads nothing for them.
    lazy: no tensors at all, only the `gen` recipe — load_scene generates each owned domain right before its commit and
    drops the tensors afterwards (C4: 64 x 8 Mi triangles per GPU would otherwise hold 32 GB of raw meshes beside the
    committed scenes)."""
    doms = []
    gx, gy, gz = grid
    for i in range(gx * gy * gz):
        ix, iy, iz = i % gx, (i // gx) % gy, i // (gx * gy)
        lo = np.array([ix * cell, iy * cell, iz * cell], np.float32)
        hi = lo + f32(cell)
        d = dict(transform=None, bounds=np.concatenate([lo, hi]).astype(np.float32), mtl=mtl, gen=(seed0 + i, tris_per_domain, lo, hi))
        if not lazy and (only is None or i in only):
            d["verts"], d["faces"], d["rgb"] = soup_domain_device(seed0 + i, tris_per_domain, lo.tolist(), hi.tolist())
        doms.append(d)
    return dict(domains=doms, lights=[(LIGHT_POINT, (0.0, 500.0, 1000.0), (1.0, 1.0, 1.0))])


def scene_bounds(scene):
    """Scene::mergeDomainBounds (reference src/render/scene.inl:296-349)."""
    b = np.stack([d["bounds"] for d in scene["domains"]]).astype(np.float32)
    return np.concatenate([b[:, :3].min(0), b[:, 3:].max(0)]).astype(np.float32)


def _normalize(v):
    v = v.astype(np.float32)
    d = f32(f32(f32(v[0] * v[0]) + f32(v[1] * v[1])) + f32(v[2] * v[2]))
    return (v * f32(f32(1.0) / np.sqrt(d, dtype=np.float32))).astype(np.float32)


def _cross(x, y):
    return np.array([f32(x[1] * y[2]) - f32(y[1] * x[2]), f32(x[2] * y[0]) - f32(y[2] * x[0]),
                     f32(x[0] * y[1]) - f32(y[0] * x[1])], np.float32)


def camera_frame(pos, lookat, up, vfov, image_w, image_h):
    """Camera::init (reference src/render/camera.h:128-160) in float32 with the same operation order."""
    pos = np.asarray(pos, np.float32); lookat = np.asarray(lookat, np.float32); up = np.asarray(up, np.float32)
    aspect = f32(f32(image_w) / f32(image_h))
    theta = f32(float(f32(vfov)) * math.pi / float(f32(180.0)))
    half_h = f32(math.tan(float(f32(theta / f32(2.0)))))
    half_w = f32(aspect * half_h)
    w = (pos - lookat).astype(np.float32)
    u = _cross(up, w)
    v = _cross(w, u)
    w, u, v = _normalize(w), _normalize(u), _normalize(v)
    center = (pos - w).astype(np.float32)
    ll = ((center - (half_w * u).astype(np.float32)).astype(np.float32) - (half_h * v).astype(np.float32)).astype(np.float32)
    wv = (f32(f32(2.0) * half_w) * u).astype(np.float32)
    hv = (f32(f32(2.0) * half_h) * v).astype(np.float32)
    return dict(pos=pos, lowerleft=ll, wvec=wv, hvec=hv, image_w=int(image_w), image_h=int(image_h))


def default_camera_pose(bounds6):
    """SprayRenderer::initCamera without --camera (reference src/render/spray_renderer.inl:87-105)."""
    b = np.asarray(bounds6, np.float32)
    ext = (b[3:] - b[:3]).astype(np.float32)
    lookat = np.array([f32(0.5 * float(f32(b[k] + b[3 + k]))) for k in range(3)], np.float32)
    length = np.sqrt(f32(f32(f32(ext[0] * ext[0]) + f32(ext[1] * ext[1])) + f32(ext[2] * ext[2])), dtype=np.float32)
    s = f32(length * f32(0.5))
    campos = np.array([lookat[0] + f32(s * f32(0.0)), lookat[1] + f32(s * f32(0.0)), lookat[2] + f32(s * f32(1.0))], np.float32)
    return campos, lookat, np.array([0.0, 1.0, 0.0], np.float32)


def image_schedule_tiles(image_w, image_h, spp, num_ranks, rank, max_samples_per_rank=1024 * 1024):
    """ImageScheduleTileList::init (reference src/render/tile.cc:317-391) with makeVerticalStripe
    (tile.cc:204-235): this rank's vertical stripe cut into horizontal blocking tiles."""
    w = max(image_w // num_ranks, 1)
    x = rank * w
    if x >= image_w:
        return []
    sw = image_w - x if (x + w > image_w or rank == num_ranks - 1) else w
    nsamples = sw * image_h * spp
    ntiles = (nsamples + max_samples_per_rank - 1) // max_samples_per_rank
    tile_h = max(image_h // ntiles, 1)
    return [(x, y, sw, min(tile_h, image_h - y)) for y in range(0, image_h, tile_h)]


def write_spray_file(scene, path, ply_name="wavelet.ply"):
    """Write a `.spray` description (tokens of reference src/io/scene_loader.cc:42-70) for a scene whose
    domains carry 'file'/'translate' (wavelets64)."""
    with open(path, "w") as f:
        f.write("# light <type> <position x y z> <intensity r g b>\n")
        for ty, pos, rad in scene["lights"]:
            if ty == LIGHT_POINT:
                f.write("light point %g %g %g %g %g %g\n" % (*pos, *rad))
            else:
                f.write("light diffuse %g %g %g\n" % tuple(rad))
        for i, d in enumerate(scene["domains"]):
            f.write("\n# domain %d\ndomain\nfile %s\nmtl diffuse 1 1 1\n" % (i, d.get("file", ply_name)))
            f.write("bound -10.000000 -10.000000 -10.000000 10.000000 9.324713 10.000000\n")
            f.write("face %d\nvertex %d\n" % (d["faces"].shape[0], d["verts"].shape[0]))
            t = d.get("translate", [0.0, 0.0, 0.0])
            f.write("translate %.6f %.6f %.6f\n" % tuple(t))


def write_ply(path, verts, faces, rgb_packed=None, binary=True):
    """PLY writer (binary little endian or ascii) in the dialect src/io/ply_loader.cc reads:
    float x,y,z [uchar red,green,blue]; list uchar int vertex_indices."""
    verts = np.asarray(verts, np.float32); faces = np.asarray(faces, np.int32)
    hdr = ["ply", "format %s 1.0" % ("binary_little_endian" if binary else "ascii"), "comment spray_b200 fixture",
           "element vertex %d" % len(verts), "property float x", "property float y", "property float z"]
    if rgb_packed is not None:
        hdr += ["property uchar red", "property uchar green", "property uchar blue"]
    hdr += ["element face %d" % len(faces), "property list uchar int vertex_indices", "end_header"]
    with open(path, "wb") as f:
        f.write(("\n".join(hdr) + "\n").encode("ascii"))
        if binary:
            if rgb_packed is not None:
                c = np.asarray(rgb_packed, np.uint32)
                vdt = np.dtype([("p", "<f4", (3,)), ("c", "u1", (3,))])
                rec = np.zeros(len(verts), vdt)
                rec["p"] = verts
                rec["c"] = np.stack([(c >> 16) & 0xff, (c >> 8) & 0xff, c & 0xff], axis=1).astype(np.uint8)
                f.write(rec.tobytes())
            else:
                f.write(verts.astype("<f4").tobytes())
            fdt = np.dtype([("n", "u1"), ("i", "<i4", (3,))])
            rec = np.zeros(len(faces), fdt)
            rec["n"] = 3; rec["i"] = faces
            f.write(rec.tobytes())
        else:
            for i, v in enumerate(verts):
                line = "%.9g %.9g %.9g" % tuple(v)
                if rgb_packed is not None:
                    c = int(rgb_packed[i])
                    line += " %d %d %d" % ((c >> 16) & 0xff, (c >> 8) & 0xff, c & 0xff)
                f.write((line + "\n").encode("ascii"))
            for t in faces:
                f.write(("3 %d %d %d\n" % tuple(t)).encode("ascii"))
