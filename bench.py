#!/usr/bin/env python3
"""bench.py — headline benchmark of the per-domain ray-tracing hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c3|c4s|c5|c4 (N > 1)]
                    [--partition image|insitu|both]

A step = one frame of the hot path on the named workload: every blocking tile of the image goes through the
fused persistent kernel (eye rays -> domain culling -> per-domain BVH8 closest hit -> shading -> shadow rays ->
frame buffer).  `value` = (primary + shadow [+ bounce] rays resolved per frame) / device time, inputs resident in
HBM; `e2e` = the same through the C ABI with a pinned HOST image (descriptors H2D, the frame's pixels D2H inside the
timed region; frames pipelined: frame k's PCIe stores overlap frame k + 1's trace, the per-frame-synchronised figure is
printed beside it).

N > 1 (torchrun, one rank per GPU), `--partition`:
  image  (headline; the reference's `spray_ooc --partition image`): every rank holds all domains and renders an
         interleaved set of row blocks; finished rows are PUSHED over NVLink into rank 0's frame buffer (CUDA IPC
         mapping, no NCCL on the frame path); e2e: every rank copies its own rows D2H into one shared pinned image.
  insitu (the reference's `spray_insitu_*`, north_star's "each GPU owns its domains' BVHs"): domains are sharded
         with the Morton partition, hit keys are merged across GPUs (64-bit min), occlusion with OR.
  both   (default at N > 1): image is the headline line, the in-situ figure is attached as "insitu".
Every N > 1 line carries "parity_check": the same frame rendered by rank 0 alone, outside the timed region.

--impl reference times the CPU oracle (the reference itself needs Embree 2 / MPI / IceT and cannot be built here)
on all host threads over the WHOLE frame of the same workload.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s (primary+shadow, device-timed)"
UNIT = "Mrays/s"
COUNTERS_JSON = os.path.join(ROOT, "profiles", "ncu_counters.json")


def workload(name, device_scene=True):
    """(scene, camera frame, shade kwargs, label, W, H).  device_scene: C3-family soups are generated on the GPU."""
    from spray_b200 import scenes
    if name == "c2":
        sc = scenes.wavelets64()
        p = scenes.WAVELETS64_CAMERA
        W = H = 2048
        cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, H)
        shade = dict(shader=0, bounces=1, pixel_samples=1, ao_samples=1, **scenes.WAVELETS64_BLINN)
        label = "C2: examples/wavelets64 (64 wavelet domains, 350720 tris), 2048x2048, PT shader, 1 point light, bounces 1, 1 spp"
    elif name == "c1":
        sc = scenes.wavelet_single()
        W = H = 1024
        pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
        cd = scenes.camera_frame(pos, lookat, up, 90.0, W, H)
        shade = dict(shader=0, bounces=1, pixel_samples=1, ao_samples=1, **scenes.WAVELETS64_BLINN)
        label = "C1: examples/wavelet.ply single domain, 1024x1024, primary+shadow"
    elif name in ("c3", "c4s"):
        tris = int(os.environ.get("SPRAY_C3_TRIS", str(1 << 20)))
        gen = os.environ.get("SPRAY_C3_GEN", "device" if device_scene else "numpy")
        sc = scenes.soup_scene_device((4, 4, 4), tris) if gen == "device" and name == "c3" else scenes.soup_scene((4, 4, 4), tris)
        W = H = 2048
        cd = scenes.camera_frame((95.0, 90.0, 85.0), (40.0, 40.0, 40.0), (0, 1, 0), 90.0, W, H)
        if name == "c3":
            shade = dict(shader=0, bounces=3, pixel_samples=1, ao_samples=1, **scenes.WAVELETS64_BLINN)
            label = "C3: synthetic 64-domain triangle soup (%s-generated), %d tris/domain, 2048x2048, bounces 3 + shadow rays" % (gen, tris)
        else:
            # C4's mechanism at a size one GPU holds: the C3 scene behind a device LRU of SPRAY_C4_CAP cache blocks
            shade = dict(shader=0, bounces=1, pixel_samples=1, ao_samples=1, **scenes.WAVELETS64_BLINN)
            label = ("C4 (reduced): 64-domain triangle soup, %d tris/domain, out-of-core with a %s-block device LRU, 2048x2048, bounces 1"
                     % (tris, os.environ.get("SPRAY_C4_CAP", "16")))
    elif name == "c4":
        # 512 domains (8 x 8 x 8 cells of extent 20), 2^23 triangles each = 4.29 G triangles: sharded over the GPUs of one box
        # (in-situ partition, 64 domains per GPU at 8 GPUs), meshes generated on the owning GPU right before their commit
        tris = int(os.environ.get("SPRAY_C4_TRIS", str(1 << 23)))
        g = int(os.environ.get("SPRAY_C4_GRID", "8"))
        sc = scenes.soup_scene_device((g, g, g), tris, lazy=True)
        W = H = 2048
        e = 20.0 * g
        cd = scenes.camera_frame((1.19 * e, 1.12 * e, 1.06 * e), (0.5 * e, 0.5 * e, 0.5 * e), (0, 1, 0), 90.0, W, H)
        shade = dict(shader=0, bounces=1, pixel_samples=1, ao_samples=1, **scenes.WAVELETS64_BLINN)
        label = "C4: synthetic %d-domain triangle soup (device-generated), %d tris/domain = %.2f G triangles, 2048x2048, bounces 1" % (g ** 3, tris, g ** 3 * tris / 1e9)
    elif name == "c5":
        # in-situ configuration: 8 wavelet domains per GPU (8 / 16 / 32 / 64 domains at 1 / 2 / 4 / 8 GPUs), 4096^2 x 4 spp,
        # multi-sample eye rays jittered with the in-situ seed (insitu_ray.h:161-165)
        world = int(os.environ.get("WORLD_SIZE", "1"))
        nd = min(8 * world, 64)
        sc = scenes.wavelets64(nd)
        W = H = int(os.environ.get("SPRAY_C5_SIZE", "4096"))
        pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
        cd = scenes.camera_frame(pos, lookat, up, 90.0, W, H)
        shade = dict(shader=0, bounces=1, pixel_samples=4, ao_samples=1, eye_seed=1, **scenes.WAVELETS64_BLINN)
        label = "C5: in situ, %d wavelet domains (8 per GPU), %dx%d, 4 spp, in-situ eye seeds" % (nd, W, H)
    else:
        raise SystemExit("unknown workload " + name)
    return sc, cd, shade, label, W, H


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for nme, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def lib_hash():
    """Build id of the libspraygpu.so this process loads (capi.build_id: sha256 of its SOURCES — nvcc's output is not
    bit-reproducible): profiles/ncu_counters.json is keyed by it, so counters of another build are never printed
    beside this one's timings."""
    from spray_b200 import capi
    return capi.build_id()


def ncu_counters(workload_name):
    """ncu --set full counters of render_tile_kernel for THIS build (tools/ncu_counters.py wrote them), else None."""
    try:
        with open(COUNTERS_JSON) as f:
            db = json.load(f)
    except Exception:
        return None, "no profiles/ncu_counters.json"
    ent = db.get(lib_hash())
    if not ent or workload_name not in ent:
        return None, "profiles/ncu_counters.json holds no capture of this build (%s...) for %s" % (lib_hash()[:12], workload_name)
    return ent[workload_name], ent[workload_name].get("source", "profiles/ncu_counters.json")


def run_cpu(sc, cd, shade, W, H, steps, warmup, tiles=None, keep=None):
    """Oracle (CPU restatement, OpenMP) on the WHOLE frame, cut into the same blocking tiles as the GPU arm (sample ids,
    and with them the seeds, are tile-relative).  Returns (Mrays/s, seconds/step, rays, cores); with `keep` (a dict) the
    last frame, its counters and the oracle scene are left in it for the parity check."""
    from oracle import pyoracle as O
    from spray_b200 import scenes
    osc = O.Scene(sc["domains"])
    cam = O.camera_from_desc(cd)
    sh, keep_alive = O.make_shade(lights=sc["lights"], **shade)
    if tiles is None:
        tiles = scenes.image_schedule_tiles(W, H, shade["pixel_samples"], 1, 0)
    times, rays = [], 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        fb, st = osc.render(cam, sh, list(tiles))
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            rays = st.radiance_rays + st.shadow_rays
    if keep is not None:
        keep.update(frame=fb, stats=st, scene=osc)
    sec = float(np.mean(times))
    return rays / sec / 1e6, sec, int(rays), O.lib().orc_num_threads()


class Timer:
    def __init__(self, stream, flush, dev, world):
        self.stream, self.flush, self.dev, self.world = stream, flush, dev, world

    def barrier(self):
        import torch
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def __call__(self, step, steps, warmup):
        import torch
        for _ in range(warmup):
            step()
        self.barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        l0 = self.dev.launch_count()
        for a, b in evs:
            self.flush.fill_(1)                           # L2 flush between timed iterations (outside the events)
            a.record(self.stream)
            step()
            b.record(self.stream)
        self.barrier()
        launches = self.dev.launch_count() - l0
        ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
        if self.world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    def fill_ms(self, reps=8):
        """Device time of one L2 fill alone (the 256 MB store between timed iterations), CUDA events, mean of `reps`."""
        import torch
        self.barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for a, b in evs:
            a.record(self.stream)
            self.flush.fill_(1)
            b.record(self.stream)
        torch.cuda.synchronize()
        return float(np.mean([a.elapsed_time(b) for a, b in evs]))

    def pipelined(self, step, join, finish, steps, warmup):
        """K frames back to back whose read-back overlaps the next frame (`step` = trace + on-device compaction of the
        frame's pixels; the PCIe stores of frame k run on the image's own stream while frame k + 1 is traced).

        ONE event pair brackets the whole sequence: opened before frame 0's L2 fill, closed after `join` has made the
        device stream wait for the last frame's stores.  The K fills inside it are taken out with a separately measured
        fill time: ms per frame = (bracket - K * fill) / K.  The stores of frame k also progress while the fill before
        frame k + 1 runs, so this can understate a pipeline whose stores are LONGER than its trace by at most one fill
        per frame; `upper` = bracket / K (fills counted as frame time) is returned beside it and printed, and the
        per-frame-synchronised figure stays in the line.  Returns (ms, upper_ms, fill_ms)."""
        import torch
        for _ in range(warmup):
            step()
        finish()
        fill = self.fill_ms()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        for _ in range(steps):
            self.flush.fill_(1)
            step()
        join()
        e1.record(self.stream)
        finish()
        self.barrier()
        total = float(e0.elapsed_time(e1))
        ms, upper = (total - steps * fill) / steps, total / steps
        if self.world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms, upper, fill], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, upper, fill = (float(x) for x in t.tolist())
        return ms, upper, fill


def single_gpu(name, dev, stream, timer, steps, warmup, with_e2e=True, keep_frame=False, key_check=False):
    """One workload on one GPU through the fused kernel.  Returns the dict the JSON line is built from."""
    import torch
    from spray_b200 import capi, scenes
    sc, cd, shade, label, W, H = workload(name)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **shade)
    tiles = scenes.image_schedule_tiles(W, H, shade["pixel_samples"], 1, 0)
    t_build0 = time.perf_counter()
    scenes_by_id = dev.load_scene(sc)
    dev.synchronize()
    build_s = time.perf_counter() - t_build0
    infos = [s_.info() for s_ in scenes_by_id.values()]
    build_dev_ms = float(sum(i.build_ms for i in infos))          # CUDA-event time of every spraygpu_scene_commit
    build_tris = int(sum(i.num_faces for i in infos))
    # steady state of a cache block that is re-loaded (out-of-core LRU miss): same buffers, no allocation
    for _ in range(2):                                    # the second commit of a shape captures its CUDA graph, later ones replay it
        for s_ in scenes_by_id.values():
            s_.commit()
    for s_ in scenes_by_id.values():
        s_.commit()
    rebuild_dev_ms = float(sum(s_.info().build_ms for s_ in scenes_by_id.values()))
    # the same re-commits ENQUEUED back to back (spraygpu_scene_commit_async), one host wait at the end: small domains
    # build concurrently on the library's pool of build streams, large ones queue on the device's stream
    dev.synchronize()
    t_b0 = time.perf_counter()
    for s_ in scenes_by_id.values():
        s_.commit(wait=False)
    for s_ in scenes_by_id.values():
        s_.info()
    rebuild_batch_ms = (time.perf_counter() - t_b0) * 1e3
    fb_d = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    kernel_events = []                                    # (start, end) around the render launch alone, per frame

    def step_dev():
        fb_d.zero_()                                      # image_.clear(), spray_renderer.inl:363
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        dev.trace_frame(cam, sh, tiles, fb_d)             # the frame's blocking tiles in one persistent launch
        b.record(stream)
        kernel_events.append((a, b))

    # ray counts (deterministic): one counted frame, also gives N_node / N_tri for the roofline
    stats_tot = {}
    fb_d.zero_()
    for t in tiles:
        st = dev.trace_tile(cam, sh, t, fb_d, stats=True, counted=True)
        for k, v in st.as_dict().items():
            stats_tot[k] = stats_tot.get(k, 0) + v
    rays = stats_tot["radiance_rays"] + stats_tot["shadow_rays"]
    ms_dev, launches = timer(step_dev, steps, warmup)
    out = dict(label=label, W=W, H=H, rays=rays, stats=stats_tot, ms_dev=ms_dev, launches=launches, tiles=tiles,
               kern_ms=float(np.mean([a.elapsed_time(b) for a, b in kernel_events[-steps:]])),
               build_s=build_s, build_dev_ms=build_dev_ms, rebuild_dev_ms=rebuild_dev_ms, rebuild_batch_ms=rebuild_batch_ms,
               build_tris=build_tris, build_domains=len(scenes_by_id),
               scene=sc, cd=cd, shade=shade)
    if with_e2e:
        # e2e = the public C-ABI frame with a pinned HOST image (HdrImage::buf): trace into the device render target,
        # then spraygpu_image_flush stores the frame's pixels into the host image over PCIe (only pixels that are or were
        # non-zero: the host image stays bit-identical to the device frame, checked below) and clears the target.
        # Camera / shade / tile descriptors go H2D as kernel arguments every frame.
        img = capi.Image(dev, None, W, H)

        def step_e2e():
            dev.trace_frame(cam, sh, tiles, img.device)
            img.flush(to_host=True, clear=True)
            dev.synchronize()

        n_e2e = max(steps // 2, 3)
        out["ms_e2e"], _ = timer(step_e2e, n_e2e, 2)
        _, pix = img.flushed_pixels()

        # the same frames back to back, frame k's PCIe stores (spraygpu_image_flush_async: compact list, a few CTAs on the
        # image's own stream) overlapping frame k + 1's trace (Timer.pipelined)
        def step_pipe():
            dev.trace_frame(cam, sh, tiles, img.device)
            img.flush_async(to_host=True)

        out["ms_e2e_pipe"], out["ms_e2e_pipe_upper"], out["ms_fill"] = timer.pipelined(step_pipe, img.join, img.wait, max(steps, 5), 3)
        img.wait()
        out["h2d"] = int(capi.C.sizeof(capi.Camera) + capi.C.sizeof(capi.Shade) + len(tiles) * 16)
        out["d2h"] = int(pix * 16 // (n_e2e + 2))
        fb_d.zero_()
        dev.trace_frame(cam, sh, tiles, fb_d)
        dev.synchronize()
        out["e2e_host_image_equals_device_frame"] = bool(np.array_equal(img.host, fb_d.cpu().numpy()))
        img.close()
        # the same with the frame copied WHOLE (64 MB at 2048^2) through spraygpu_trace_tile_async, round 1's e2e path
        fb_h = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory()
        fb_h_np = fb_h.numpy()

        def step_full():
            for t in tiles:
                dev.trace_tile_async(cam, sh, t, fb_h_np)
            dev.synchronize()

        out["ms_e2e_full"], _ = timer(step_full, 3, 1)
    if key_check:
        try:
            out["key_check"] = soup_key_check(dev, cam, sc, W, H)
        except Exception as e:
            out["key_check"] = {"error": repr(e)[:300]}
    if keep_frame:
        # parity material for the bench line (outside every timed region): the frame itself and the closest-hit keys of
        # all its eye rays through the batched seam
        fb_d.zero_()
        dev.trace_frame(cam, sh, tiles, fb_d)
        dev.synchronize()
        out["frame"] = fb_d.cpu().numpy()
        if shade["pixel_samples"] == 1:
            n = W * H
            rays_d = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
            dev.gen_eye_rays(cam, (0, 0, W, H), 1, rays_d)
            keys_d = torch.zeros(n, dtype=torch.int64, device="cuda")
            dev.intersect_domains(rays_d, n, keys_d, None)
            dev.synchronize()
            out["eye_rays_host"] = rays_d.cpu().numpy()
            out["eye_keys"] = keys_d.cpu().numpy().view(np.uint64)
            del rays_d, keys_d
    for s_ in list(scenes_by_id.values()):
        s_.close()
    dev.scenes = [s_ for s_ in dev.scenes if s_.h]
    return out


def soup_key_check(dev, cam, sc, W, H, window=320):
    """The benched C3 scene against the oracle: the closest-hit keys of a window of >= 100 k eye rays from the device
    (batched seam, the same BVHs the timed frames traverse) against the CPU checker, which REGENERATES the triangles of
    every domain those rays can enter from the same counter-based function (bit-identical meshes: tests/test_gpu_sizes.py)
    and builds its own BVH over them.  Domains no sampled ray enters get a one-triangle stand-in (never tested)."""
    import torch
    from oracle import pyoracle as O
    from spray_b200 import scenes
    x0, y0 = (W - window) // 2, (H - window) // 2
    n = window * window
    rays_d = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
    dev.gen_eye_rays(cam, (x0, y0, window, window), 1, rays_d)
    keys_d = torch.zeros(n, dtype=torch.int64, device="cuda")
    dev.intersect_domains(rays_d, n, keys_d, None)
    dev.synchronize()
    rays = rays_d.cpu().numpy(); keys = keys_d.cpu().numpy().view(np.uint64)
    o, d = rays[:, 0:3].astype(np.float64), rays[:, 4:7].astype(np.float64)
    inv = 1.0 / np.where(d == 0.0, 1e-300, d)
    t0 = time.perf_counter()
    doms_o, rebuilt = [], 0
    for dm in sc["domains"]:
        b = np.asarray(dm["bounds"], np.float64)
        ta, tb = (b[:3] - 1e-3 - o) * inv, (b[3:] + 1e-3 - o) * inv          # slab test, box grown: a superset of the entered domains
        enter = (np.minimum(ta, tb).max(1) <= np.maximum(ta, tb).min(1)) & (np.maximum(ta, tb).min(1) > 0)
        seed, ntris, lo, hi = dm["gen"]
        if enter.any():
            v, f, c = O.gen_soup(0x5EED0000 + int(seed), int(ntris), lo, hi, scenes.soup_h(ntris, lo, hi))
            rebuilt += 1
        else:
            v = np.tile(0.5 * (np.asarray(lo, np.float32) + np.asarray(hi, np.float32)), (3, 1)); f = np.array([[0, 1, 2]], np.uint32); c = np.zeros(3, np.uint32)
        doms_o.append(dict(verts=v, faces=f, rgb=c, transform=None, bounds=np.asarray(dm["bounds"], np.float32), mtl=dm["mtl"]))
    osc = O.Scene(doms_o)
    ko, _ = osc.intersect(rays)
    return {"against": "oracle O2 over meshes regenerated on the CPU from the same Philox function", "eye_rays": int(n),
            "window": [x0, y0, window, window], "domains_regenerated": rebuilt, "keys_equal": bool(np.array_equal(keys, ko)),
            "keys_differing": int((keys != ko).sum()), "hit_fraction": float((keys != O.KEY_MISS).mean()),
            "cpu_seconds": time.perf_counter() - t0}


def parity_check_of(r, kept):
    """GPU frame of the benched workload against the oracle frame the cpu_baseline leg rendered anyway (same blocking
    tiles): RMSE / max |diff| over RGB (north_star: RMSE <= 1e-4), equal ray counters, and — at 1 spp — the closest-hit
    key of EVERY eye ray (stride 1) bit-equal to the oracle's."""
    g, o = r["frame"], kept["frame"]
    d = g[..., :3].astype(np.float64) - o[..., :3].astype(np.float64)
    st_g, st_o = r["stats"], kept["stats"]
    pc = {"against": "oracle O2 (CPU restatement), whole %dx%d frame" % (r["W"], r["H"]),
          "rmse": float(np.sqrt(np.mean(d * d))), "max_abs": float(np.abs(d).max()), "tolerance_rmse": 1e-4,
          "ray_counts_equal": bool(int(st_g["eye_rays"]) == int(st_o.eye_rays) and int(st_g["radiance_rays"]) == int(st_o.radiance_rays)
                                   and int(st_g["shadow_rays"]) == int(st_o.shadow_rays))}
    if "eye_keys" in r:
        ko, _ = kept["scene"].intersect(r["eye_rays_host"])
        pc["keys_equal_stride1"] = bool(np.array_equal(r["eye_keys"], ko))
        pc["keys_compared"] = int(len(ko))
    pc["ok"] = bool(pc["rmse"] <= 1e-4 and pc["ray_counts_equal"] and pc.get("keys_equal_stride1", True))
    return pc


def roofline_of(r, name, steps):
    """algorithmic bytes = 16 B/pixel frame buffer + 80 B/node + 48 B/triangle + 48 B attribute record per
    hit, N_* counted by the counters build of the same kernel on the same rays; achieved = that / the kernel's
    CUDA-event duration; peak = measured HBM copy bandwidth.  Beside it the ncu counters of THIS build (if captured):
    they say what actually bounds the kernel."""
    peak, peak_src = measured_peak_hbm()
    st = r["stats"]
    nn, nt, nh, nrays = st["nodes_visited"], st["tris_tested"], st["hits"], r["rays"]
    bytes_frame = 16 * r["W"] * r["H"] + 80 * nn + 48 * nt + 48 * nh
    nl = max(int(r["launches"]) // max(steps, 1), 1)      # render launches per frame (1: spraygpu_trace_frame)
    per_launch = bytes_frame / nl
    dur_s = r["kern_ms"] * 1e-3 / nl
    achieved = per_launch / dur_s / 1e9
    ctr, src = ncu_counters(name)
    traffic = (ctr["dram_bytes_read"] + ctr["dram_bytes_write"]) if ctr else None
    if ctr is None:
        bound = "hbm"
        note = "algorithmic-bytes rate / HBM peak; no ncu capture of this build, so the limiter is not measured here"
    elif traffic < 0.25 * per_launch:
        bound = "latency"
        note = ("DRAM traffic is %.1f %% of the algorithmic bytes (node / triangle fetches are served by L1 / L2): the kernel is "
                "latency- and divergence-bound, frac is an algorithmic-bytes rate over the HBM peak, not a DRAM utilisation"
                % (100.0 * traffic / per_launch))
    else:
        bound = "hbm"
        note = "DRAM traffic is %.0f %% of the algorithmic bytes" % (100.0 * traffic / per_launch)
    rf = {"bound": bound, "bound_note": note, "kernel": "render_tile_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
          "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src, "traffic_source": src,
          "algorithmic_bytes_per_launch": per_launch, "launch_ms": dur_s * 1e3,
          "nodes_per_ray": nn / nrays, "tris_per_ray": nt / nrays, "bytes_per_ray": bytes_frame / nrays,
          "domain_tests_per_ray": st["domain_tests"] / nrays}
    for k in ("l2_bytes", "l1_throughput_pct", "l2_throughput_pct", "dram_throughput_pct", "issue_slots_busy_pct", "lanes_per_inst",
              "l1_hit_pct", "l2_hit_pct", "registers", "achieved_warps_per_sm", "ncu_duration_ms"):
        rf[k] = ctr.get(k) if ctr else None
    return rf


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("SPRAY_WORKLOAD", "c2"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-workloads", action="store_true", help="skip the C3 figure attached to the default C2 line")
    ap.add_argument("--partition", default=os.environ.get("SPRAY_PARTITION", "both"), choices=["image", "insitu", "both"],
                    help="N > 1: image = the reference's --partition image (all domains on every rank, interleaved row "
                         "blocks, rows pushed to rank 0 over NVLink); insitu = --partition insitu (domains sharded, hit keys "
                         "merged with a 64-bit min across GPUs); both = headline image, insitu reported beside it")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        # torchrun exports OMP_NUM_THREADS=1; the CPU arm uses every host core (set before libgomp loads)
        os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
        # The reference's own CPU implementation cannot be built (Embree 2.x, MPI, IceT, OpenGL absent);
        # this arm times the oracle port on all host threads.  Rank 0 only.
        if rank != 0:
            return
        sc, cd, shade, label, W, H = workload(args.workload, device_scene=False)
        from spray_b200 import scenes as _scenes
        tiles_ref = _scenes.image_schedule_tiles(W, H, shade["pixel_samples"], 1, 0)
        kept_ref = {}
        v, sec, rays, cores = run_cpu(sc, cd, shade, W, H, max(args.steps, 1), min(args.warmup, 1), tiles=tiles_ref, keep=kept_ref)
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic camera rays on the reference's example scene",
                # the same keys as the GPU arm's config (the driver compares the two dicts' keys)
                "config": {"workload": label, "tiles_per_frame": len(tiles_ref), "l2_flush_between_steps": False,
                           "rays_per_frame": rays, "eye_rays": int(kept_ref["stats"].eye_rays), "shadow_rays": int(kept_ref["stats"].shadow_rays),
                           "hits": None, "parallelism": "%d host threads (OpenMP), all domains resident" % cores,
                           "build_s_all_domains": None, "lib_build_id": None},
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                 "sample": "the whole %dx%d frame (%d rays/step); CPU restatement (oracle O2), NOT Embree" % (W, H, rays)},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    from spray_b200 import capi, scenes
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libspraygpu has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    stream = torch.cuda.current_stream()
    dev = capi.Device(local_rank, stream=stream.cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    timer = Timer(stream, flush, dev, world)
    sampler = ClockSampler(local_rank)

    if world > 1:
        from spray_b200 import multigpu
        line = multigpu.bench_multi(args, dev, stream, timer, sampler, rank, world, local_rank, workload, METRIC, UNIT)
        if rank == 0:
            comp = line.pop("_composite", None)
            # the N-rank composite against the ORACLE's frame (GPU-vs-GPU identity alone would not notice an error both share):
            # C1 / C2 only — the CPU restatement needs seconds for those, hours for the soups
            if comp is not None and args.workload in ("c1", "c2") and not args.no_cpu_baseline and isinstance(line.get("parity_check"), dict):
                try:
                    try:                                  # torchrun pins OMP_NUM_THREADS=1 and libgomp is already initialised
                        import ctypes
                        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(max((os.cpu_count() or 1) // world, 1))
                    except Exception:
                        pass
                    sc, cd, shade, label, W, H = workload(args.workload, device_scene=False)
                    kept = {}
                    _, sec, _, cores = run_cpu(sc, cd, shade, W, H, 1, 0, keep=kept)
                    d = comp[..., :3].astype(np.float64) - kept["frame"][..., :3].astype(np.float64)
                    line["parity_check"]["oracle_rmse"] = float(np.sqrt(np.mean(d * d)))
                    line["parity_check"]["oracle_max_abs"] = float(np.abs(d).max())
                    line["parity_check"]["oracle_note"] = "composite of %d ranks vs oracle O2 (CPU restatement), whole frame, %d threads, %.1f s" % (world, cores, sec)
                    line["summary"]["oracle_rmse"] = line["parity_check"]["oracle_rmse"]
                except Exception as e:
                    line["parity_check"]["oracle_error"] = repr(e)[:200]
            line["summary"] = line.pop("summary")         # stays the last key
            print(json.dumps(line))
        dist.destroy_process_group()
        dev.close()
        return

    if args.workload == "c4s":
        line = bench_c4s(args, dev, stream, timer, sampler)
        print(json.dumps(line))
        dev.close()
        return

    sampler.start()
    r = single_gpu(args.workload, dev, stream, timer, args.steps, args.warmup, keep_frame=not args.no_cpu_baseline)
    clocks = sampler.stop()
    value = r["rays"] / (r["ms_dev"] * 1e-3) / 1e6
    e2e_sync_v = r["rays"] / (r["ms_e2e"] * 1e-3) / 1e6
    e2e_ms = min(r["ms_e2e"], r["ms_e2e_pipe"])
    e2e_v = r["rays"] / (e2e_ms * 1e-3) / 1e6
    st = r["stats"]
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_dev"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic camera rays on the reference's example scene (random-free)",
            "config": {"workload": r["label"], "tiles_per_frame": len(r["tiles"]), "l2_flush_between_steps": True,
                       "rays_per_frame": int(r["rays"]), "eye_rays": int(st["eye_rays"]), "shadow_rays": int(st["shadow_rays"]),
                       "hits": int(st["hits"]),
                       "parallelism": "1 GPU, all domains resident", "build_s_all_domains": r["build_s"],
                       "lib_build_id": lib_hash()[:16]},
            "clocks": clocks, "gpu_launches": int(r["launches"]),
            "e2e": {"value": e2e_v, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": r["h2d"],
                    "d2h_bytes_per_step": r["d2h"],
                    "mode": "pipelined frames: spraygpu_image_flush_async compacts frame k's pixels on the device and a few CTAs on the image's "
                            "own stream store them into the pinned host image while frame k + 1 is traced; one event pair around all frames up to the "
                            "moment the last frame's stores have landed, minus the separately timed L2 fills" if r["ms_e2e_pipe"] <= r["ms_e2e"] else "one frame at a time",
                    "per_frame_sync": {"value": e2e_sync_v, "ms_per_step": r["ms_e2e"],
                                       "note": "trace, flush, host synchronize per frame (no overlap between frames); L2 fill outside the events"},
                    "pipelined_ms_per_step": r["ms_e2e_pipe"],
                    "pipelined_ms_per_step_upper": r["ms_e2e_pipe_upper"], "l2_fill_ms": r["ms_fill"],
                    "pipelined_note": "upper = the same bracket with the L2 fills counted as frame time (frame k's stores also progress during the fill before frame k + 1)",
                    "d2h_note": "spraygpu_image_flush: pixels that are or were non-zero are stored into the pinned host image "
                                "(full frame = %d B); host image == device frame: %s" % (r["W"] * r["H"] * 16, r["e2e_host_image_equals_device_frame"]),
                    "full_copy": {"value": r["rays"] / (r["ms_e2e_full"] * 1e-3) / 1e6, "ms_per_step": r["ms_e2e_full"],
                                  "d2h_bytes_per_step": r["W"] * r["H"] * 16, "api": "spraygpu_trace_tile_async, whole rows copied"}}}
    line["roofline"] = roofline_of(r, args.workload, args.steps)
    # per-domain acceleration-structure build (= rtcCommit): device time of the commits alone; the wall-clock figure
    # in config also holds the H2D upload of the meshes from pageable host memory
    line["build"] = {"triangles": r["build_tris"], "first_commit_device_ms": r["build_dev_ms"],
                     "recommit_device_ms": r["rebuild_dev_ms"],
                     "mtris_per_s": r["build_tris"] / max(r["rebuild_dev_ms"], 1e-9) / 1e3,
                     "domains": r["build_domains"],
                     "recommit_all_async_wall_ms": r["rebuild_batch_ms"],
                     "ms_per_domain_amortised": r["rebuild_batch_ms"] / max(r["build_domains"], 1),
                     "mtris_per_s_async": r["build_tris"] / max(r["rebuild_batch_ms"], 1e-9) / 1e3,
                     "algorithmic_gb_per_s": r["build_tris"] * 300.0 / max(r["rebuild_dev_ms"], 1e-9) / 1e6,
                     "note": "first commit includes cudaMalloc of the domain's buffers; mtris_per_s is the re-commit"}
    if args.workload == "c2" and not args.no_extra_workloads:
        # C3 (64 x 1 Mi triangles, bounces 3: the configuration large enough to leave L2) in the same record
        try:
            r3 = single_gpu("c3", dev, stream, timer, max(min(args.steps, 10), 3), 3, with_e2e=False, key_check=not args.no_cpu_baseline)
            rf3 = roofline_of(r3, "c3", max(min(args.steps, 10), 3))
            line["workloads"] = {"c3": {"workload": r3["label"], "value": r3["rays"] / (r3["ms_dev"] * 1e-3) / 1e6, "unit": UNIT,
                                        "ms_per_step": r3["ms_dev"], "rays_per_frame": int(r3["rays"]), "roofline": rf3,
                                        "keys_equal": (r3.get("key_check") or {}).get("keys_equal"), "key_check": r3.get("key_check"),
                                        "frac": rf3["frac"], "nodes_per_ray": rf3["nodes_per_ray"], "tris_per_ray": rf3["tris_per_ray"],
                                        "build": {"triangles": r3["build_tris"], "recommit_device_ms": r3["rebuild_dev_ms"],
                                                  "mtris_per_s": r3["build_tris"] / max(r3["rebuild_dev_ms"], 1e-9) / 1e3,
                                                  "recommit_all_async_wall_ms": r3["rebuild_batch_ms"]}}}
        except Exception as e:                              # never lose the headline line to the attachment
            line["workloads"] = {"c3": {"error": repr(e)[:300]}}
    if not args.no_cpu_baseline:
        sc, cd, shade, label, W, H = workload(args.workload, device_scene=False) if args.workload != "c2" else (r["scene"], r["cd"], r["shade"], r["label"], r["W"], r["H"])
        nsteps = 3 if args.workload in ("c1", "c2") else 1
        kept = {}
        v, sec, crays, cores = run_cpu(sc, cd, shade, W, H, nsteps, 1 if nsteps > 1 else 0, tiles=r["tiles"], keep=kept)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": "the whole frame (%d rays), %d steps; CPU restatement (oracle O2), NOT Embree" % (crays, nsteps)}
        if "frame" in r:
            try:
                line["parity_check"] = parity_check_of(r, kept)
            except Exception as e:
                line["parity_check"] = {"error": repr(e)[:300]}
    print(json.dumps(line))
    dev.close()


def bench_c4s(args, dev, stream, timer, sampler):
    """C4's mechanism at a size one GPU holds: out-of-core sweep, meshes wait in PINNED host memory, the device LRU
    uploads + commits on every miss (spraygpu_trace_tile_ooc)."""
    import torch
    from spray_b200 import capi, scenes
    sc, cd, shade, label, W, H = workload("c4s")
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **shade)
    cap = int(os.environ.get("SPRAY_C4_CAP", "16"))
    # one blocking tile for the whole frame (the reference's --max-samples-per-rank, config.cc:66, raised to W*H):
    # every sweep loads each entered domain once per FRAME instead of once per 1 Mi-sample tile
    tiles = [(0, 0, W, H)]
    rt = torch.cuda.cudart()
    for d in sc["domains"]:
        d["verts"] = np.ascontiguousarray(d["verts"], np.float32)
        d["faces"] = np.ascontiguousarray(d["faces"], np.uint32)
        d["rgb"] = np.ascontiguousarray(d["rgb"], np.uint32)
        for k in ("verts", "faces", "rgb"):
            rt.cudaHostRegister(d[k].ctypes.data, d[k].nbytes, 0)     # page-lock in place
    cache = capi.LruDomainCache(dev, sc, cap)
    fb_d = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    fb_h = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory()
    ooc_stats = {}

    def step_dev():
        fb_d.zero_()
        for t in tiles:
            dev.trace_tile_ooc(cam, sh, t, fb_d, cap, cache.load)

    def step_e2e():
        step_dev()
        fb_h.copy_(fb_d, non_blocking=True)
        dev.synchronize()

    fb_d.zero_()
    for t in tiles:
        st = dev.trace_tile_ooc(cam, sh, t, fb_d, cap, cache.load, stats=True)
        for k, v in st.as_dict().items():
            ooc_stats[k] = ooc_stats.get(k, 0) + v
    misses_per_frame = cache.misses
    rays = ooc_stats["radiance_rays"] + ooc_stats["shadow_rays"]
    sampler.start()
    ms_dev, launches = timer(step_dev, args.steps, args.warmup)
    clocks = sampler.stop()
    ms_e2e, _ = timer(step_e2e, max(args.steps // 2, 3), 2)
    mesh_bytes = int(sum(d["verts"].nbytes + d["faces"].nbytes + d["rgb"].nbytes for d in sc["domains"]) // len(sc["domains"]))
    return {"metric": METRIC, "value": rays / (ms_dev * 1e-3) / 1e6, "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic triangle soup",
            "config": {"workload": label, "tiles_per_frame": len(tiles), "l2_flush_between_steps": True, "rays_per_frame": int(rays),
                       "parallelism": "1 GPU, out-of-core: %d cache blocks for %d domains" % (cap, len(sc["domains"]))},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": rays / (ms_e2e * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(misses_per_frame) * mesh_bytes, "d2h_bytes_per_step": int(W * H * 16)},
            "ooc": {"cache_blocks": cap, "mesh_bytes_per_domain": mesh_bytes, "domains": len(sc["domains"]),
                    "misses_first_frame": int(misses_per_frame), "misses_total": int(cache.misses), "hits_total": int(cache.hits),
                    "note": "every miss = H2D of the domain's mesh from pinned memory + spraygpu_scene_commit"}}


if __name__ == "__main__":
    main()
