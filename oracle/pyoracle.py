"""ctypes binding of the CPU oracle (oracle/spray_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing under spray_b200/ or host/ imports this module.  parity unpinned (see
spray_oracle.h): the reference has no tests or golden vectors and cannot be built here.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libspray_oracle.so")

KEY_MISS = 0xFFFFFFFFFFFFFFFF
FLAG_NEAR_TIE = 1
FLAG_EDGE = 2


def build(force=False):
    """Compile the oracle with gcc (oracle/Makefile).  No-op when the .so is newer than its sources."""
    srcs = [os.path.join(_HERE, f) for f in ("spray_oracle.c", "spray_oracle.h", "Makefile")]
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= max(os.path.getmtime(s) for s in srcs)):
        return _LIB_PATH
    env = dict(os.environ)
    env.pop("CC", None)
    subprocess.run(["make", "-C", _HERE], check=True, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return _LIB_PATH


class Camera(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("lowerleft", C.c_float * 3), ("wvec", C.c_float * 3),
                ("hvec", C.c_float * 3), ("image_w", C.c_int), ("image_h", C.c_int)]


class Tile(C.Structure):
    _fields_ = [("x", C.c_int), ("y", C.c_int), ("w", C.c_int), ("h", C.c_int)]


class Light(C.Structure):
    _fields_ = [("type", C.c_int), ("position", C.c_float * 3), ("radiance", C.c_float * 3)]


class Shade(C.Structure):
    _fields_ = [("shader", C.c_int), ("bounces", C.c_int), ("pixel_samples", C.c_int), ("ao_samples", C.c_int),
                ("ks", C.c_float * 3), ("shininess", C.c_float), ("nlights", C.c_int), ("lights", C.POINTER(Light)),
                ("eye_seed", C.c_int)]


class Stats(C.Structure):
    _fields_ = [("eye_rays", C.c_uint64), ("radiance_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("domain_tests_radiance", C.c_uint64), ("domain_tests_shadow", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        build()
    L = C.CDLL(_LIB_PATH)
    f32p, u32p, i32p, u64p, u8p, f64p = (C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_int32),
                                          C.POINTER(C.c_uint64), C.POINTER(C.c_uint8), C.POINTER(C.c_double))
    vp = C.c_void_p
    L.orc_rng_init1.restype = C.c_uint32; L.orc_rng_init1.argtypes = [C.c_int]
    L.orc_rng_init2.restype = C.c_uint32; L.orc_rng_init2.argtypes = [C.c_int, C.c_int]
    L.orc_rng_get_float.restype = C.c_float; L.orc_rng_get_float.argtypes = [u32p]
    L.orc_morton.restype = C.c_uint32; L.orc_morton.argtypes = [C.c_float] * 3
    L.orc_intersect_aabb.restype = C.c_int
    L.orc_intersect_aabb.argtypes = [f32p, f32p, f32p, C.c_float, C.c_float, f32p, f32p]
    L.orc_interp_color.restype = C.c_uint32
    L.orc_interp_color.argtypes = [C.c_uint32] * 3 + [C.c_float] * 2
    L.orc_unpack_color.argtypes = [C.c_uint32, f32p]
    L.orc_camera_init.argtypes = [f32p, f32p, f32p, C.c_float, C.c_int, C.c_int, C.POINTER(Camera)]
    L.orc_default_camera_pose.argtypes = [f32p] * 4
    L.orc_camera_ray.argtypes = [C.POINTER(Camera), C.c_float, C.c_float, f32p, f32p]
    L.orc_image_schedule_tiles.restype = C.c_int
    L.orc_image_schedule_tiles.argtypes = [C.c_int64] * 4 + [C.c_int, C.c_int64, C.POINTER(Tile), C.c_int]
    L.orc_partition.argtypes = [C.c_int, f32p, f32p, C.c_int, i32p, u32p]
    L.orc_cosine_hemisphere.argtypes = [C.c_float, C.c_float, f32p, f32p, f32p]
    L.orc_blinn_phong.argtypes = [C.c_float, f32p, f32p, C.c_float, f32p, f32p, f32p, f32p, f32p]
    L.orc_ppm_quantize.argtypes = [f32p, C.c_int, u32p]
    L.orc_tri_test.argtypes = [f32p] * 5 + [C.c_float, i32p, f32p, f32p, f32p]
    L.orc_lru_sequence.argtypes = [C.c_int, C.c_int, i32p, C.c_int, i32p, i32p]
    L.orc_scene_create.restype = vp; L.orc_scene_create.argtypes = [C.c_int]
    L.orc_scene_destroy.argtypes = [vp]
    L.orc_scene_set_domain.argtypes = [vp, C.c_int, f32p, C.c_size_t, u32p, C.c_size_t, u32p, f32p, f32p, C.c_int, f32p]
    L.orc_scene_commit.argtypes = [vp]
    L.orc_scene_bounds.argtypes = [vp, f32p]
    L.orc_scene_num_verts.restype = C.c_size_t; L.orc_scene_num_verts.argtypes = [vp, C.c_int]
    L.orc_scene_get_normals.argtypes = [vp, C.c_int, f32p]
    L.orc_scene_get_verts.argtypes = [vp, C.c_int, f32p]
    L.orc_gen_eye_rays.argtypes = [C.POINTER(Camera), Tile, C.c_int, vp]
    L.orc_gen_eye_rays_seed.argtypes = [C.POINTER(Camera), Tile, C.c_int, C.c_int, vp]
    L.orc_domain_lists.argtypes = [vp, vp, C.c_size_t, C.c_int, i32p, i32p, f32p]
    L.orc_intersect.argtypes = [vp, vp, C.c_size_t, u64p, f32p]
    L.orc_intersect_brute.argtypes = [vp, vp, C.c_size_t, u64p, f32p]
    L.orc_intersect_domain.argtypes = [vp, C.c_int, vp, C.c_size_t, u64p, f32p]
    L.orc_occluded.argtypes = [vp, vp, C.c_size_t, u8p]
    L.orc_occluded_domain.argtypes = [vp, C.c_int, vp, C.c_size_t, u8p]
    L.orc_intersect_f64.argtypes = [vp, vp, C.c_size_t, C.c_int, i32p, i32p, f64p, u8p]
    L.orc_hit_attributes.argtypes = [vp, u64p, f32p, C.c_size_t, u32p, f32p]
    L.orc_render_tile.argtypes = [vp, C.POINTER(Camera), C.POINTER(Shade), Tile, f32p, C.POINTER(Stats)]
    L.orc_num_threads.restype = C.c_int
    L.orc_philox4x32_10.argtypes = [u32p, u32p, u32p]
    L.orc_gen_soup.argtypes = [C.c_uint32, C.c_size_t, f32p, f32p, C.c_float, f32p, u32p, u32p]
    _lib = L
    return L


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def make_camera(campos, lookat, up, vfov, w, h):
    cam = Camera()
    lib().orc_camera_init(_p(_f32(campos), C.c_float), _p(_f32(lookat), C.c_float), _p(_f32(up), C.c_float),
                          C.c_float(vfov), w, h, C.byref(cam))
    return cam


def camera_from_desc(desc):
    """desc: dict with pos/lowerleft/wvec/hvec/image_w/image_h (spray_b200.scenes.camera_frame)."""
    cam = Camera()
    for k in ("pos", "lowerleft", "wvec", "hvec"):
        for i in range(3):
            getattr(cam, k)[i] = float(np.float32(desc[k][i]))
    cam.image_w, cam.image_h = int(desc["image_w"]), int(desc["image_h"])
    return cam


def default_camera_pose(bounds6):
    campos, lookat, up = np.zeros(3, np.float32), np.zeros(3, np.float32), np.zeros(3, np.float32)
    lib().orc_default_camera_pose(_p(_f32(bounds6), C.c_float), _p(campos, C.c_float), _p(lookat, C.c_float),
                                  _p(up, C.c_float))
    return campos, lookat, up


def image_schedule_tiles(w, h, spp, nranks, rank, max_samples):
    buf = (Tile * 4096)()
    n = lib().orc_image_schedule_tiles(w, h, spp, nranks, rank, max_samples, buf, 4096)
    return [(buf[i].x, buf[i].y, buf[i].w, buf[i].h) for i in range(n)]


def make_shade(shader=0, bounces=1, pixel_samples=1, ao_samples=1, ks=(0.4, 0.4, 0.4), shininess=10.0, lights=(),
               eye_seed=0):
    """lights: sequence of (type, position3, radiance3).  eye_seed 1 = the in-situ tracers' (pixid, s) jitter seed.
    Returns (Shade, keepalive)."""
    arr = (Light * max(1, len(lights)))()
    for i, (ty, pos, rad) in enumerate(lights):
        arr[i].type = ty
        for k in range(3):
            arr[i].position[k] = float(pos[k]); arr[i].radiance[k] = float(rad[k])
    sh = Shade()
    sh.shader, sh.bounces, sh.pixel_samples, sh.ao_samples = shader, bounces, pixel_samples, ao_samples
    for k in range(3):
        sh.ks[k] = float(ks[k])
    sh.shininess = float(shininess)
    sh.nlights = len(lights)
    sh.eye_seed = int(eye_seed)
    sh.lights = C.cast(arr, C.POINTER(Light))
    return sh, arr


class Scene:
    """Oracle scene built from a spray_b200.scenes.SceneDesc-like object (list of domain dicts)."""

    def __init__(self, domains):
        L = lib()
        self.nd = len(domains)
        self.h = L.orc_scene_create(self.nd)
        self._keep = []
        for i, d in enumerate(domains):
            verts = _f32(d["verts"]).reshape(-1, 3)
            faces = np.ascontiguousarray(d["faces"], dtype=np.uint32).reshape(-1, 3)
            rgb = d.get("rgb")
            rgb = None if rgb is None else np.ascontiguousarray(rgb, dtype=np.uint32)
            m = d.get("transform")
            m = None if m is None else _f32(m).reshape(16)
            bounds = _f32(d["bounds"]).reshape(6)
            mtl_type, mtl = d.get("mtl", (0, (1.0, 1.0, 1.0)))
            mtl = _f32(mtl)
            L.orc_scene_set_domain(self.h, i, _p(verts, C.c_float), verts.shape[0], _p(faces, C.c_uint32),
                                   faces.shape[0], None if rgb is None else _p(rgb, C.c_uint32),
                                   None if m is None else _p(m, C.c_float), _p(bounds, C.c_float), int(mtl_type),
                                   _p(mtl, C.c_float))
        L.orc_scene_commit(self.h)

    def __del__(self):
        try:
            if self.h:
                lib().orc_scene_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def bounds(self):
        out = np.zeros(6, np.float32)
        lib().orc_scene_bounds(self.h, _p(out, C.c_float))
        return out

    def normals(self, dom):
        nv = lib().orc_scene_num_verts(self.h, dom)
        out = np.zeros((nv, 3), np.float32)
        lib().orc_scene_get_normals(self.h, dom, _p(out, C.c_float))
        return out

    def verts(self, dom):
        nv = lib().orc_scene_num_verts(self.h, dom)
        out = np.zeros((nv, 3), np.float32)
        lib().orc_scene_get_verts(self.h, dom, _p(out, C.c_float))
        return out

    @staticmethod
    def _rays(rays):
        r = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        return r, r.ctypes.data_as(C.c_void_p), r.shape[0]

    def domain_lists(self, rays, cap=32):
        r, rp, n = self._rays(rays)
        counts = np.zeros(n, np.int32); ids = np.full((n, cap), -1, np.int32); tm = np.zeros((n, cap), np.float32)
        lib().orc_domain_lists(self.h, rp, n, cap, _p(counts, C.c_int32), _p(ids, C.c_int32), _p(tm, C.c_float))
        return counts, ids, tm

    def intersect(self, rays, brute=False):
        r, rp, n = self._rays(rays)
        keys = np.zeros(n, np.uint64); uv = np.zeros((n, 2), np.float32)
        fn = lib().orc_intersect_brute if brute else lib().orc_intersect
        fn(self.h, rp, n, _p(keys, C.c_uint64), _p(uv, C.c_float))
        return keys, uv

    def intersect_domain(self, dom, rays):
        r, rp, n = self._rays(rays)
        keys = np.zeros(n, np.uint64); uv = np.zeros((n, 2), np.float32)
        lib().orc_intersect_domain(self.h, dom, rp, n, _p(keys, C.c_uint64), _p(uv, C.c_float))
        return keys, uv

    def occluded(self, rays):
        r, rp, n = self._rays(rays)
        occ = np.zeros(n, np.uint8)
        lib().orc_occluded(self.h, rp, n, _p(occ, C.c_uint8))
        return occ

    def occluded_domain(self, dom, rays):
        r, rp, n = self._rays(rays)
        occ = np.zeros(n, np.uint8)
        lib().orc_occluded_domain(self.h, dom, rp, n, _p(occ, C.c_uint8))
        return occ

    def intersect_f64(self, rays, use_bvh=True):
        r, rp, n = self._rays(rays)
        dom = np.zeros(n, np.int32); prim = np.zeros(n, np.int32); t = np.zeros(n, np.float64)
        flags = np.zeros(n, np.uint8)
        lib().orc_intersect_f64(self.h, rp, n, int(use_bvh), _p(dom, C.c_int32), _p(prim, C.c_int32),
                                _p(t, C.c_double), _p(flags, C.c_uint8))
        return dom, prim, t, flags

    def hit_attributes(self, keys, uv):
        keys = np.ascontiguousarray(keys, np.uint64); uv = _f32(uv)
        n = keys.shape[0]
        color = np.zeros(n, np.uint32); ns = np.zeros((n, 3), np.float32)
        lib().orc_hit_attributes(self.h, _p(keys, C.c_uint64), _p(uv, C.c_float), n, _p(color, C.c_uint32),
                                 _p(ns, C.c_float))
        return color, ns

    def render(self, cam, shade, tiles, fb=None):
        """tiles: list of (x,y,w,h).  Returns (fb[h,w,4], Stats)."""
        if fb is None:
            fb = np.zeros((cam.image_h, cam.image_w, 4), np.float32)
        st = Stats()
        for (x, y, w, h) in tiles:
            lib().orc_render_tile(self.h, C.byref(cam), C.byref(shade), Tile(x, y, w, h), _p(fb, C.c_float),
                                  C.byref(st))
        return fb, st


def gen_eye_rays(cam, tile, spp=1, eye_seed=0):
    x, y, w, h = tile
    rays = np.zeros((w * h * spp, 8), np.float32)
    lib().orc_gen_eye_rays_seed(C.byref(cam), Tile(x, y, w, h), spp, int(eye_seed), rays.ctypes.data_as(C.c_void_p))
    return rays


def key_fields(keys):
    keys = np.asarray(keys, np.uint64)
    t = (keys >> np.uint64(32)).astype(np.uint32).view(np.float32)
    dom = ((keys >> np.uint64(23)) & np.uint64(0x1FF)).astype(np.int32)
    prim = (keys & np.uint64(0x7FFFFF)).astype(np.int32)
    return t, dom, prim


def philox4x32_10(ctr, key):
    c = np.asarray(ctr, np.uint32); k = np.asarray(key, np.uint32); out = np.zeros(4, np.uint32)
    lib().orc_philox4x32_10(_p(c, C.c_uint32), _p(k, C.c_uint32), _p(out, C.c_uint32))
    return out


def soup_h(ntris, lo, hi):
    """Mean triangle spacing of a soup cell, the float both generators are handed: extent / ntris^(1/3)."""
    ext = np.asarray(hi, np.float32) - np.asarray(lo, np.float32)
    return np.float32(float(ext.max()) / float(ntris) ** (1.0 / 3.0))


def gen_soup(key, ntris, lo, hi, h=None):
    """orc_gen_soup: (verts f32[3n,3], faces u32[n,3], rgb u32[3n]) of the counter-based soup domain."""
    lo = np.ascontiguousarray(lo, np.float32); hi = np.ascontiguousarray(hi, np.float32)
    if h is None:
        h = soup_h(ntris, lo, hi)
    xyz = np.zeros((3 * ntris, 3), np.float32); idx = np.zeros((ntris, 3), np.uint32); rgb = np.zeros(3 * ntris, np.uint32)
    lib().orc_gen_soup(int(key) & 0xFFFFFFFF, ntris, _p(lo, C.c_float), _p(hi, C.c_float), float(h), _p(xyz, C.c_float),
                       _p(idx, C.c_uint32), _p(rgb, C.c_uint32))
    return xyz, idx, rgb
