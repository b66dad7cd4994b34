"""ctypes binding of libspraygpu.so (include/spray_gpu.h) — the reference-facing C ABI.

Loading the library never needs a GPU (so `-m "not gpu"` tests can check the exported symbols); every
compute entry point needs one and raises SprayGpuError otherwise.  There is no CPU fallback.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# SPRAYGPU_LIB selects an experimental build variant of the SAME sources (tools/variants.sh)
LIB_PATH = os.environ.get("SPRAYGPU_LIB") or os.path.join(_HERE, "libspraygpu.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "spray_gpu.h")
CSRC_DIR = os.path.join(_HERE, "csrc")

KEY_MISS = 0xFFFFFFFFFFFFFFFF
MAX_LIGHTS = 8


class SprayGpuError(RuntimeError):
    pass


def build(force=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... (spray_b200/csrc/Makefile).  Cross-compiles
    without a GPU."""
    csrc = os.path.join(_HERE, "csrc")
    srcs = [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cu", ".cuh", "Makefile"))] + [HEADER_PATH]
    if (not force and os.path.exists(LIB_PATH)
            and os.path.getmtime(LIB_PATH) >= max(os.path.getmtime(s) for s in srcs)):
        return LIB_PATH
    r = subprocess.run(["make", "-C", csrc], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise SprayGpuError("building libspraygpu.so failed:\n" + r.stdout[-4000:])
    return LIB_PATH


class Ray(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("ox", "oy", "oz", "tmin", "dx", "dy", "dz", "tmax")]


class Material(C.Structure):
    _fields_ = [("type", C.c_int32), ("params", C.c_float * 3)]


class Light(C.Structure):
    _fields_ = [("type", C.c_int32), ("position", C.c_float * 3), ("radiance", C.c_float * 3)]


class Camera(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("lowerleft", C.c_float * 3), ("wvec", C.c_float * 3),
                ("hvec", C.c_float * 3), ("image_w", C.c_int32), ("image_h", C.c_int32)]


class Tile(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32), ("w", C.c_int32), ("h", C.c_int32)]


class Shade(C.Structure):
    _fields_ = [("shader", C.c_int32), ("bounces", C.c_int32), ("pixel_samples", C.c_int32),
                ("ao_samples", C.c_int32), ("ks", C.c_float * 3), ("shininess", C.c_float), ("nlights", C.c_int32),
                ("lights", Light * MAX_LIGHTS), ("eye_seed", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("eye_rays", "radiance_rays", "shadow_rays", "domain_tests",
                                           "nodes_visited", "tris_tested", "hits")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class SceneInfo(C.Structure):
    _fields_ = [("num_vertices", C.c_uint64), ("num_faces", C.c_uint64), ("num_nodes", C.c_uint64),
                ("node_bytes", C.c_uint64), ("tri_bytes", C.c_uint64), ("attr_bytes", C.c_uint64),
                ("bounds", C.c_float * 6), ("build_ms", C.c_double), ("bvh_depth", C.c_uint32), ("reserved", C.c_uint32)]


_vp = C.c_void_p
_PROTOS = {
    "spraygpu_last_error": (C.c_char_p, []),
    "spraygpu_version": (C.c_char_p, []),
    "spraygpu_device_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "spraygpu_device_destroy": (C.c_int, [_vp]),
    "spraygpu_device_set_stream": (C.c_int, [_vp, _vp]),
    "spraygpu_device_synchronize": (C.c_int, [_vp]),
    "spraygpu_device_launch_count": (C.c_uint64, [_vp]),
    "spraygpu_scene_create": (C.c_int, [_vp, C.POINTER(_vp)]),
    "spraygpu_scene_destroy": (C.c_int, [_vp]),
    "spraygpu_scene_set_mesh": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp]),
    "spraygpu_scene_set_transform": (C.c_int, [_vp, _vp]),
    "spraygpu_scene_commit": (C.c_int, [_vp]),
    "spraygpu_scene_commit_async": (C.c_int, [_vp]),
    "spraygpu_scene_get_bvh": (C.c_int, [_vp, _vp, _vp]),
    "spraygpu_scene_get_info": (C.c_int, [_vp, C.POINTER(SceneInfo)]),
    "spraygpu_scene_get_normals": (C.c_int, [_vp, _vp]),
    "spraygpu_scene_get_vertices": (C.c_int, [_vp, _vp]),
    "spraygpu_compute_normals_static": (C.c_int, [_vp, C.c_size_t, _vp, C.c_size_t, _vp, _vp]),
    "spraygpu_gen_soup": (C.c_int, [C.c_uint32, C.c_size_t, _vp, _vp, C.c_float, _vp, _vp, _vp, _vp]),
    "spraygpu_intersect": (C.c_int, [_vp, C.c_int, _vp, C.c_size_t, _vp, _vp]),
    "spraygpu_occluded": (C.c_int, [_vp, _vp, C.c_size_t, _vp]),
    "spraygpu_domains_set": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "spraygpu_domain_bind": (C.c_int, [_vp, C.c_int, _vp]),
    "spraygpu_domain_lists": (C.c_int, [_vp, _vp, C.c_size_t, C.c_int, _vp, _vp, _vp]),
    "spraygpu_intersect_domains": (C.c_int, [_vp, _vp, C.c_size_t, _vp, _vp, C.c_int]),
    "spraygpu_occluded_domains": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_int]),
    "spraygpu_hit_attributes": (C.c_int, [_vp, _vp, _vp, C.c_size_t, _vp, _vp]),
    "spraygpu_gen_eye_rays": (C.c_int, [_vp, C.POINTER(Camera), Tile, C.c_int, _vp]),
    "spraygpu_gen_eye_rays_insitu": (C.c_int, [_vp, C.POINTER(Camera), Tile, C.c_int, _vp]),
    "spraygpu_trace_tile": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(Shade), Tile, _vp, C.POINTER(Stats)]),
    "spraygpu_trace_tile_counted": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(Shade), Tile, _vp, C.POINTER(Stats)]),
    "spraygpu_trace_tile_async": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(Shade), Tile, _vp]),
    "spraygpu_trace_frame": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(Shade), _vp, C.c_int, _vp, C.POINTER(Stats)]),
    "spraygpu_trace_frame_rects": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(Shade), _vp, _vp, C.c_int, _vp, C.POINTER(Stats)]),
    "spraygpu_trace_tile_ooc": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(Shade), Tile, _vp, C.c_int, _vp, _vp, C.POINTER(Stats)]),
    # multi-GPU (NCCL resolved at run time)
    "spraygpu_comm_unique_id": (C.c_int, [C.c_char_p]),
    "spraygpu_comm_create": (C.c_int, [_vp, C.c_int, C.c_int, C.c_char_p, C.POINTER(_vp)]),
    "spraygpu_comm_destroy": (C.c_int, [_vp]),
    "spraygpu_partition_domains": (C.c_int, [C.c_int, _vp, _vp, C.c_int, C.c_int, _vp, _vp]),
    "spraygpu_trace_tile_insitu": (C.c_int, [_vp, _vp, C.POINTER(Camera), C.POINTER(Shade), Tile, _vp, C.POINTER(Stats)]),
    "spraygpu_trace_tile_insitu_p2p": (C.c_int, [_vp, _vp, C.POINTER(Camera), C.POINTER(Shade), Tile, _vp, C.c_int, C.POINTER(Stats)]),
    "spraygpu_trace_tile_insitu_fwd": (C.c_int, [_vp, _vp, C.POINTER(Camera), C.POINTER(Shade), Tile, _vp, C.c_int, _vp, C.POINTER(Stats)]),
    "spraygpu_image_reduce": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int]),
    "spraygpu_image_gather_rows": (C.c_int, [_vp, _vp, _vp, C.c_int, _vp, C.c_int]),
    "spraygpu_image_gather_blocks": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, _vp, _vp, C.c_int]),
    "spraygpu_image_create": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_char_p, C.POINTER(_vp)]),
    "spraygpu_image_destroy": (C.c_int, [_vp]),
    "spraygpu_image_device": (_vp, [_vp]),
    "spraygpu_image_host": (_vp, [_vp]),
    "spraygpu_image_root": (_vp, [_vp]),
    "spraygpu_image_set_rows": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "spraygpu_image_flush": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int]),
    "spraygpu_image_flush_async": (C.c_int, [_vp, C.c_int, C.c_int]),
    "spraygpu_image_wait": (C.c_int, [_vp]),
    "spraygpu_image_join": (C.c_int, [_vp]),
    "spraygpu_image_flushed_pixels": (C.c_int, [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "spraygpu_image_barrier": (C.c_int, [_vp]),
    "spraygpu_image_read": (C.c_int, [_vp, C.c_int, _vp]),
    "spraygpu_allreduce_keys_min": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "spraygpu_allreduce_occ_max": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
}
UNIQUE_ID_BYTES = 128

_lib = None


def lib():
    """Load libspraygpu.so.  Raises (loudly) if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SprayGpuError("libspraygpu.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'` "
                            "or `make -C spray_b200/csrc`); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(L, name)  # AttributeError if a declared symbol is missing
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def build_id():
    """Identifies the library build for profiles/ncu_counters.json: sha256 over the SOURCES libspraygpu.so is made from
    (csrc/*.cu, *.cuh, Makefile, include/spray_gpu.h).  nvcc's output is not bit-reproducible (two clean builds of the
    same tree differ), so a hash of the binary would stop matching as soon as anybody rebuilds.  A variant library
    loaded through SPRAYGPU_LIB (other -D flags, same sources) is identified by its own bytes instead."""
    import glob
    import hashlib
    h = hashlib.sha256()
    if os.environ.get("SPRAYGPU_LIB"):
        with open(LIB_PATH, "rb") as f:
            h.update(f.read())
        return "variant-" + h.hexdigest()
    files = sorted(glob.glob(os.path.join(CSRC_DIR, "*.cu")) + glob.glob(os.path.join(CSRC_DIR, "*.cuh")) +
                   [os.path.join(CSRC_DIR, "Makefile"), HEADER_PATH])
    for fn in files:
        h.update(os.path.basename(fn).encode())
        with open(fn, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def declared_symbols():
    """Entry points declared in include/spray_gpu.h."""
    import re
    with open(HEADER_PATH) as f:
        return sorted(set(re.findall(r"SPRAYGPU_API\s+[\w\s\*]+?\b(spraygpu_\w+)\s*\(", f.read())))


def _check(rc):
    if rc != 0:
        raise SprayGpuError("libspraygpu error %d: %s" % (rc, lib().spraygpu_last_error().decode()))


def _ptr(a):
    """numpy array -> host pointer; torch tensor -> its data_ptr (device or host); None -> NULL."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data
    if hasattr(a, "data_ptr"):
        assert a.is_contiguous()
        return a.data_ptr()
    return a


def gen_soup(key, ntris, lo, hi, h, xyz, idx=None, rgb=None, cuda_stream=None):
    """spraygpu_gen_soup: the counter-based triangle soup of configs C3 / C4 written into CUDA tensors (xyz f32[9n],
    idx i32[3n] or None, rgb i32[3n] or None) on the current device."""
    lo = np.ascontiguousarray(lo, np.float32); hi = np.ascontiguousarray(hi, np.float32)
    _check(lib().spraygpu_gen_soup(int(key) & 0xFFFFFFFF, int(ntris), lo.ctypes.data, hi.ctypes.data, float(np.float32(h)),
                                   _ptr(xyz), _ptr(idx), _ptr(rgb), _vp(cuda_stream) if cuda_stream else None))


STATIC_MAX_VERTS = 1 << 17
STATIC_MAX_FACES = 1 << 18


def compute_normals_static(xyz, idx, normals, cuda_stream=None):
    """spraygpu_compute_normals_static: per-vertex normals of a mesh that already lives on the GPU, written into
    `normals`, with the library's built-in scratch array as work space — the call allocates nothing, for hosts (an
    in-situ simulation that owns all of HBM) where a plug-in's cudaMalloc is not an option.  xyz / idx / normals are
    CUDA tensors (or raw device addresses together with explicit counts via lib()); needs no Device."""
    nv, nf = xyz.numel() // 3, idx.numel() // 3
    assert xyz.element_size() == 4 and idx.element_size() == 4 and normals.numel() == 3 * nv
    _check(lib().spraygpu_compute_normals_static(_ptr(xyz), nv, _ptr(idx), nf, _ptr(normals),
                                                 _vp(cuda_stream) if cuda_stream else None))


def make_camera(desc):
    cam = Camera()
    for k in ("pos", "lowerleft", "wvec", "hvec"):
        for i in range(3):
            getattr(cam, k)[i] = float(np.float32(desc[k][i]))
    cam.image_w, cam.image_h = int(desc["image_w"]), int(desc["image_h"])
    return cam


EYE_SEED_OOC, EYE_SEED_INSITU = 0, 1


def make_shade(shader=0, bounces=1, pixel_samples=1, ao_samples=1, ks=(0.4, 0.4, 0.4), shininess=10.0, lights=(),
               eye_seed=EYE_SEED_OOC):
    sh = Shade()
    sh.shader, sh.bounces, sh.pixel_samples, sh.ao_samples = shader, bounces, pixel_samples, ao_samples
    sh.eye_seed = int(eye_seed)
    for k in range(3):
        sh.ks[k] = float(ks[k])
    sh.shininess = float(shininess)
    sh.nlights = len(lights)
    for i, (ty, pos, rad) in enumerate(lights):
        sh.lights[i].type = ty
        for k in range(3):
            sh.lights[i].position[k] = float(pos[k]); sh.lights[i].radiance[k] = float(rad[k])
    return sh


class Device:
    def __init__(self, cuda_device=0, stream=None):
        self.h = _vp()
        _check(lib().spraygpu_device_create(int(cuda_device), C.byref(self.h)))
        self.scenes = []
        if stream is not None:
            self.set_stream(stream)

    def set_stream(self, cuda_stream_ptr):
        _check(lib().spraygpu_device_set_stream(self.h, _vp(cuda_stream_ptr)))

    def synchronize(self):
        _check(lib().spraygpu_device_synchronize(self.h))

    def launch_count(self):
        return int(lib().spraygpu_device_launch_count(self.h))

    def close(self):
        if self.h:
            for s in self.scenes:
                s.close()
            self.scenes = []
            lib().spraygpu_device_destroy(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- domains
    def set_domains(self, bounds6, materials=None):
        b = np.ascontiguousarray(bounds6, np.float32).reshape(-1, 6)
        n = b.shape[0]
        m = None
        if materials is not None:
            m = (Material * n)()
            for i, (ty, params) in enumerate(materials):
                m[i].type = int(ty)
                for k in range(3):
                    m[i].params[k] = float(params[k])
        _check(lib().spraygpu_domains_set(self.h, n, b.ctypes.data, C.cast(m, _vp) if m is not None else None))
        self.ndomains = n

    def bind(self, domain_id, scene):
        _check(lib().spraygpu_domain_bind(self.h, int(domain_id), scene.h if scene is not None else None))

    def new_scene(self):
        s = Scene(self)
        self.scenes.append(s)
        return s

    def load_scene(self, scene_desc, owned=None):
        """Upload + commit every domain of a spray_b200.scenes description and bind it.  `owned`: iterable of
        domain ids resident on this GPU (default all) — the in-situ partition."""
        doms = scene_desc["domains"]
        self.set_domains(np.stack([d["bounds"] for d in doms]), [d.get("mtl", (0, (1, 1, 1))) for d in doms])
        owned = set(range(len(doms))) if owned is None else set(int(i) for i in owned)
        out = {}
        for i, d in enumerate(doms):
            if i not in owned:
                continue
            verts, faces, rgb = d.get("verts"), d.get("faces"), d.get("rgb")
            if verts is None and d.get("gen") is not None:
                from . import scenes as _scenes            # lazily generated on the device (scenes.soup_scene_device(lazy=True))
                seed, ntris, lo, hi = d["gen"]
                verts, faces, rgb = _scenes.soup_domain_device(seed, ntris, np.asarray(lo).tolist(), np.asarray(hi).tolist())
            if verts is None:
                raise SprayGpuError("load_scene: domain %d is owned but carries no mesh" % i)
            s = self.new_scene()
            s.set_mesh(verts, faces, rgb)
            s.set_transform(d.get("transform"))
            s.commit(wait=False)                           # the builds queue up on the device; the first trace waits
            self.bind(i, s)
            out[i] = s
            del verts, faces, rgb
        return out

    # ---- batched queries (numpy host arrays or torch device tensors)
    def domain_lists(self, rays, n, cap, counts, ids, tmins):
        _check(lib().spraygpu_domain_lists(self.h, _ptr(rays), n, cap, _ptr(counts), _ptr(ids), _ptr(tmins)))

    def intersect_domains(self, rays, n, keys, uv=None, merge=False):
        _check(lib().spraygpu_intersect_domains(self.h, _ptr(rays), n, _ptr(keys), _ptr(uv), int(merge)))

    def occluded_domains(self, rays, n, occ, merge=False):
        _check(lib().spraygpu_occluded_domains(self.h, _ptr(rays), n, _ptr(occ), int(merge)))

    def hit_attributes(self, rays, keys, n, color, ns3):
        _check(lib().spraygpu_hit_attributes(self.h, _ptr(rays), _ptr(keys), n, _ptr(color), _ptr(ns3)))

    def gen_eye_rays(self, cam, tile, spp, rays, insitu=False):
        fn = lib().spraygpu_gen_eye_rays_insitu if insitu else lib().spraygpu_gen_eye_rays
        _check(fn(self.h, C.byref(cam), Tile(*tile), spp, _ptr(rays)))

    def trace_tile_async(self, cam, shade, tile, fb):
        """Host frame buffer: returns once the tile is enqueued; rows are valid after synchronize()."""
        _check(lib().spraygpu_trace_tile_async(self.h, C.byref(cam), C.byref(shade), Tile(*tile), _ptr(fb)))

    def trace_tile(self, cam, shade, tile, fb, stats=False, counted=False):
        st = Stats() if stats else None
        fn = lib().spraygpu_trace_tile_counted if counted else lib().spraygpu_trace_tile
        _check(fn(self.h, C.byref(cam), C.byref(shade), Tile(*tile), _ptr(fb), C.byref(st) if st is not None else None))
        return st


    def trace_frame(self, cam, shade, tiles, fb, stats=False):
        """Every blocking tile of a frame in one persistent launch (device fb) / overlapped per-tile copies (host fb)."""
        st = Stats() if stats else None
        arr = (Tile * len(tiles))(*[Tile(*t) for t in tiles])
        _check(lib().spraygpu_trace_frame(self.h, C.byref(cam), C.byref(shade), C.cast(arr, _vp), len(tiles), _ptr(fb),
                                          C.byref(st) if st is not None else None))
        return st

    def trace_frame_rects(self, cam, shade, rects, blocking_tiles, fb, stats=False):
        """Rectangles of blocking tiles in one persistent launch; samples keep their blocking tile's samid."""
        st = Stats() if stats else None
        assert len(rects) == len(blocking_tiles)
        ra = (Tile * max(len(rects), 1))(*[Tile(*t) for t in rects])
        pa = (Tile * max(len(rects), 1))(*[Tile(*t) for t in blocking_tiles])
        _check(lib().spraygpu_trace_frame_rects(self.h, C.byref(cam), C.byref(shade), C.cast(ra, _vp), C.cast(pa, _vp), len(rects),
                                                _ptr(fb), C.byref(st) if st is not None else None))
        return st

    def trace_tile_ooc(self, cam, shade, tile, fb, max_resident, load, stats=False):
        """Out-of-core tile: `load(domain_id) -> 0` makes a domain resident (Scene::load)."""
        st = Stats() if stats else None

        def _cb(user, domain_id):
            try:
                return int(load(int(domain_id)) or 0)
            except Exception as e:   # never unwind through the C frame
                self._cb_error = e
                return 1
        cb = LOAD_FN(_cb)
        self._cb_error = None
        rc = lib().spraygpu_trace_tile_ooc(self.h, C.byref(cam), C.byref(shade), Tile(*tile), _ptr(fb), int(max_resident),
                                           C.cast(cb, _vp), None, C.byref(st) if st is not None else None)
        if self._cb_error is not None:
            raise self._cb_error
        _check(rc)
        return st


LOAD_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int)


class LruDomainCache:
    """Device-side domain cache for the out-of-core path: the reference's LruCache (src/render/lru_cache.cc:65-171)
    driving Scene::load (src/render/scene.inl:135-187) — a miss uploads the domain's mesh into the cache block's
    scene, commits it (BVH build) and binds it; the evicted domain is unbound.  Harness for tests / bench; the
    C++ mirror is host/spray/scene.h."""

    def __init__(self, dev, scene_desc, capacity):
        self.dev, self.doms, self.capacity = dev, scene_desc["domains"], int(capacity)
        dev.set_domains(np.stack([d["bounds"] for d in self.doms]), [d.get("mtl", (0, (1, 1, 1))) for d in self.doms])
        self.blocks = [dev.new_scene() for _ in range(self.capacity)]
        self.block_domain = [-1] * self.capacity
        self.lru = []            # block indices, least recently used first
        self.where = {}          # domain id -> block
        self.hits = self.misses = 0

    def load(self, domain_id):
        b = self.where.get(domain_id)
        if b is not None:
            self.hits += 1
            self.lru.remove(b); self.lru.append(b)
            return 0
        self.misses += 1
        if len(self.where) < self.capacity:
            b = len(self.where)
        else:
            b = self.lru.pop(0)
            old = self.block_domain[b]
            self.dev.bind(old, None)
            del self.where[old]
        d = self.doms[domain_id]
        s = self.blocks[b]
        s.set_mesh(d["verts"], d["faces"], d.get("rgb"))
        s.set_transform(d.get("transform"))
        s.commit()
        self.dev.bind(domain_id, s)
        self.block_domain[b] = domain_id
        self.where[domain_id] = b
        self.lru.append(b)
        return 0


def partition_domains(world_bounds, scene_bounds, nranks, round_robin=False, return_codes=False):
    """InsituPartition::partition (reference src/render/data_partition.h:59-137) — host function of the C ABI."""
    wb = np.ascontiguousarray(world_bounds, np.float32).reshape(-1, 6)
    sb = np.ascontiguousarray(scene_bounds, np.float32).reshape(6)
    d2r = np.zeros(len(wb), np.int32)
    codes = np.zeros(len(wb), np.uint32)
    _check(lib().spraygpu_partition_domains(len(wb), wb.ctypes.data, sb.ctypes.data, int(nranks), int(round_robin),
                                            d2r.ctypes.data, codes.ctypes.data))
    return (d2r, codes) if return_codes else d2r


def comm_unique_id():
    """ncclGetUniqueId; rank 0 calls it and the application broadcasts the bytes."""
    buf = C.create_string_buffer(UNIQUE_ID_BYTES)
    _check(lib().spraygpu_comm_unique_id(buf))
    return buf.raw


class Comm:
    """One rank of the NCCL communicator that replaces the reference's MPI world."""

    def __init__(self, dev, nranks, rank, unique_id):
        self.dev, self.nranks, self.rank = dev, int(nranks), int(rank)
        self.h = _vp()
        assert len(unique_id) == UNIQUE_ID_BYTES
        _check(lib().spraygpu_comm_create(dev.h, self.nranks, self.rank, unique_id, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().spraygpu_comm_destroy(self.h)
            self.h = _vp()

    def trace_tile_insitu(self, cam, shade, tile, fb, stats=False):
        st = Stats() if stats else None
        _check(lib().spraygpu_trace_tile_insitu(self.dev.h, self.h, C.byref(cam), C.byref(shade), Tile(*tile), _ptr(fb),
                                                C.byref(st) if st is not None else None))
        return st

    def trace_tile_insitu_p2p(self, cam, shade, tile, fb, rows_per_block, stats=False):
        """The in-situ tile over peer memory (insitu2.cu): fb receives this rank's HOME rows of the final image."""
        st = Stats() if stats else None
        _check(lib().spraygpu_trace_tile_insitu_p2p(self.dev.h, self.h, C.byref(cam), C.byref(shade), Tile(*tile), _ptr(fb),
                                                    int(rows_per_block), C.byref(st) if st is not None else None))
        return st

    def trace_tile_insitu_fwd(self, cam, shade, tile, fb, rows_per_block, domain_owner, stats=False):
        """The in-situ tile with ray forwarding (insitu3.cu): rays travel to the owners of the domains they enter, every
        sample is shaded on its home rank; fb receives this rank's HOME rows of the final image."""
        st = Stats() if stats else None
        own = np.ascontiguousarray(domain_owner, np.int32)
        _check(lib().spraygpu_trace_tile_insitu_fwd(self.dev.h, self.h, C.byref(cam), C.byref(shade), Tile(*tile), _ptr(fb),
                                                    int(rows_per_block), own.ctypes.data, C.byref(st) if st is not None else None))
        return st

    def image_reduce(self, fb, npixels, root=0):
        _check(lib().spraygpu_image_reduce(self.dev.h, self.h, _ptr(fb), int(npixels), int(root)))

    def image_gather_rows(self, fb, image_w, row_begin, root=0):
        rb = np.ascontiguousarray(row_begin, np.int32)
        assert len(rb) == self.nranks + 1
        _check(lib().spraygpu_image_gather_rows(self.dev.h, self.h, _ptr(fb), int(image_w), rb.ctypes.data, int(root)))

    def image_gather_blocks(self, fb, image_w, block_row_begin, block_owner, root=0):
        rb = np.ascontiguousarray(block_row_begin, np.int32)
        ow = np.ascontiguousarray(block_owner, np.int32)
        assert len(rb) == len(ow) + 1
        _check(lib().spraygpu_image_gather_blocks(self.dev.h, self.h, _ptr(fb), int(image_w), len(ow), rb.ctypes.data,
                                                  ow.ctypes.data, int(root)))

    def allreduce_keys_min(self, keys, n):
        _check(lib().spraygpu_allreduce_keys_min(self.dev.h, self.h, _ptr(keys), int(n)))

    def allreduce_occ_max(self, occ, n):
        _check(lib().spraygpu_allreduce_occ_max(self.dev.h, self.h, _ptr(occ), int(n)))


class Image:
    """The frame image (HdrImage on the GPU path): device render target + pinned host pixels (+ the root's composite
    at N > 1).  `device` / `root` are raw device addresses (pass them as fb); `host` is a numpy view."""

    def __init__(self, dev, comm, w, h, shm_name=None):
        self.dev, self.comm, self.w, self.h = dev, comm, int(w), int(h)
        self.h_ = _vp()
        _check(lib().spraygpu_image_create(dev.h, comm.h if comm is not None else None, self.w, self.h,
                                           shm_name.encode() if shm_name else None, C.byref(self.h_)))
        self.device = lib().spraygpu_image_device(self.h_)
        self.root = lib().spraygpu_image_root(self.h_)
        hp = lib().spraygpu_image_host(self.h_)
        self.host = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_float)), shape=(self.h, self.w, 4))

    def set_rows(self, row_begin, owner=None):
        rb = np.ascontiguousarray(row_begin, np.int32)
        ow = None if owner is None else np.ascontiguousarray(owner, np.int32)
        _check(lib().spraygpu_image_set_rows(self.h_, len(rb) - 1, rb.ctypes.data, None if ow is None else ow.ctypes.data))

    def flush(self, to_root=False, to_host=False, clear=True):
        _check(lib().spraygpu_image_flush(self.h_, int(to_root), int(to_host), int(clear)))

    def flush_async(self, to_root=False, to_host=False):
        """Flush + clear with the PCIe stores overlapped with the next frame (compaction on the device's stream, a few
        CTAs on the image's own stream store the list into the host image)."""
        _check(lib().spraygpu_image_flush_async(self.h_, int(to_root), int(to_host)))

    def wait(self):
        _check(lib().spraygpu_image_wait(self.h_))

    def join(self):
        """The device stream waits for every pending flush (no host wait)."""
        _check(lib().spraygpu_image_join(self.h_))

    def flushed_pixels(self):
        a, b = C.c_uint64(), C.c_uint64()
        _check(lib().spraygpu_image_flushed_pixels(self.h_, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def barrier(self):
        _check(lib().spraygpu_image_barrier(self.h_))

    def read_device(self, which="device"):
        """D2H copy of the render target / root composite (tests, parity checks)."""
        out = np.zeros((self.h, self.w, 4), np.float32)
        _check(lib().spraygpu_image_read(self.h_, 0 if which == "device" else 1, out.ctypes.data))
        return out

    def close(self):
        if self.h_:
            lib().spraygpu_image_destroy(self.h_)
            self.h_ = _vp()


class Scene:
    def __init__(self, dev):
        self.dev = dev
        self.h = _vp()
        _check(lib().spraygpu_scene_create(dev.h, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().spraygpu_scene_destroy(self.h)
            self.h = _vp()

    def set_mesh(self, verts, faces, rgb=None):
        if hasattr(verts, "data_ptr"):
            # torch tensors (CUDA or pinned host): passed by pointer, copied device-side by the library
            assert verts.is_contiguous() and faces.is_contiguous() and (rgb is None or rgb.is_contiguous())
            assert verts.element_size() == 4 and faces.element_size() == 4
            nv, nf = verts.numel() // 3, faces.numel() // 3
            _check(lib().spraygpu_scene_set_mesh(self.h, verts.data_ptr(), nv, faces.data_ptr(), nf,
                                                 None if rgb is None else rgb.data_ptr()))
            self.nv, self.nf = nv, nf
            return
        v = np.ascontiguousarray(verts, np.float32).reshape(-1, 3)
        f = np.ascontiguousarray(faces, np.uint32).reshape(-1, 3)
        c = None if rgb is None else np.ascontiguousarray(rgb, np.uint32)
        _check(lib().spraygpu_scene_set_mesh(self.h, v.ctypes.data, v.shape[0], f.ctypes.data, f.shape[0],
                                             None if c is None else c.ctypes.data))
        self.nv, self.nf = v.shape[0], f.shape[0]

    def set_transform(self, m):
        if m is None:
            _check(lib().spraygpu_scene_set_transform(self.h, None))
        else:
            m = np.ascontiguousarray(m, np.float32).reshape(16)
            _check(lib().spraygpu_scene_set_transform(self.h, m.ctypes.data))

    def commit(self, wait=True):
        """rtcCommit: build the BVH8.  wait=False only enqueues the build (spraygpu_scene_commit_async)."""
        _check(lib().spraygpu_scene_commit(self.h) if wait else lib().spraygpu_scene_commit_async(self.h))

    def bvh(self):
        """(nodes [num_nodes, 20] uint32, tris [num_faces, 12] float32): parity dump of the committed BVH8."""
        inf = self.info()
        nodes = np.zeros((inf.num_nodes, 20), np.uint32)
        tris = np.zeros((inf.num_faces, 12), np.float32)
        _check(lib().spraygpu_scene_get_bvh(self.h, nodes.ctypes.data, tris.ctypes.data))
        return nodes, tris

    def info(self):
        inf = SceneInfo()
        _check(lib().spraygpu_scene_get_info(self.h, C.byref(inf)))
        return inf

    def normals(self):
        out = np.zeros((self.nv, 3), np.float32)
        _check(lib().spraygpu_scene_get_normals(self.h, out.ctypes.data))
        return out

    def vertices(self):
        out = np.zeros((self.nv, 3), np.float32)
        _check(lib().spraygpu_scene_get_vertices(self.h, out.ctypes.data))
        return out

    def intersect(self, domain_id, rays, n, keys, uv=None):
        _check(lib().spraygpu_intersect(self.h, int(domain_id), _ptr(rays), n, _ptr(keys), _ptr(uv)))

    def occluded(self, rays, n, occ):
        _check(lib().spraygpu_occluded(self.h, _ptr(rays), n, _ptr(occ)))
