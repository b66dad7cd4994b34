#!/usr/bin/env python3
"""bench.py — headline benchmark of the per-domain ray-tracing hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c3|c4s|c5]

A step = one frame of the hot path on the named workload: every blocking tile of the image goes through
spraygpu_trace_tile (eye rays -> domain culling -> per-domain BVH8 closest hit -> shading -> shadow rays
-> frame buffer).  `value` = (primary + shadow [+ bounce] rays resolved per frame) / device time, inputs
resident in HBM; `e2e` = the same through the C ABI with a pinned HOST frame buffer (D2H inside the timed
region).  N > 1 (torchrun, one rank per GPU): domains are sharded with the reference's Morton in-situ
partition and hits are resolved with an NCCL min-all-reduce on the packed 64-bit keys
(spray_b200/multigpu.py).

--impl reference times the CPU oracle (the reference itself needs Embree 2 / MPI / IceT and cannot be
built here) on all host threads over a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# dram__bytes_read.sum + dram__bytes_write.sum of ONE render launch (= one frame) from the committed ncu --set full
# capture of this command (bench.py cannot run under ncu itself); refreshed with every profile round.
NCU_TRAFFIC = {"c2": 18517504 + 2016768, "c3": 675356160 + 45847040}
NCU_TRAFFIC_SOURCE = "ncu --set full, profiles/r01y_{c2,c3}_render_tile_ncu_full_details.csv (one launch = one frame)"

METRIC = "Mrays/s (primary+shadow, device-timed)"
UNIT = "Mrays/s"


def workload(name):
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
    elif name == "c3":
        tris = int(os.environ.get("SPRAY_C3_TRIS", str(1 << 20)))
        sc = scenes.soup_scene((4, 4, 4), tris)
        p = scenes.WAVELETS64_CAMERA
        W = H = 2048
        cd = scenes.camera_frame((95.0, 90.0, 85.0), (40.0, 40.0, 40.0), (0, 1, 0), 90.0, W, H)
        shade = dict(shader=0, bounces=3, pixel_samples=1, ao_samples=1, **scenes.WAVELETS64_BLINN)
        label = "C3: synthetic 64-domain triangle soup, %d tris/domain, 2048x2048, bounces 3 + shadow rays" % tris
    elif name == "c4s":
        # C4's mechanism at a size one GPU holds: the C3 scene behind a device LRU of SPRAY_C4_CAP cache blocks
        # (default 16 of 64 domains) -> spraygpu_trace_tile_ooc; every sweep re-uploads and re-builds what it evicted
        tris = int(os.environ.get("SPRAY_C3_TRIS", str(1 << 20)))
        sc = scenes.soup_scene((4, 4, 4), tris)
        W = H = 2048
        cd = scenes.camera_frame((95.0, 90.0, 85.0), (40.0, 40.0, 40.0), (0, 1, 0), 90.0, W, H)
        shade = dict(shader=0, bounces=1, pixel_samples=1, ao_samples=1, **scenes.WAVELETS64_BLINN)
        label = ("C4 (reduced): 64-domain triangle soup, %d tris/domain, out-of-core with a %s-block device LRU, 2048x2048, bounces 1"
                 % (tris, os.environ.get("SPRAY_C4_CAP", "16")))
    elif name == "c5":
        # in-situ configuration: 8 wavelet domains per GPU (8 / 16 / 32 / 64 domains at 1 / 2 / 4 / 8 GPUs), 4096^2 x 4 spp
        world = int(os.environ.get("WORLD_SIZE", "1"))
        nd = min(8 * world, 64)
        sc = scenes.wavelets64(nd)
        W = H = int(os.environ.get("SPRAY_C5_SIZE", "4096"))
        pos, lookat, up = scenes.default_camera_pose(scenes.scene_bounds(sc))
        cd = scenes.camera_frame(pos, lookat, up, 90.0, W, H)
        shade = dict(shader=0, bounces=1, pixel_samples=4, ao_samples=1, **scenes.WAVELETS64_BLINN)
        label = "C5: in situ, %d wavelet domains (8 per GPU), %dx%d, 4 spp" % (nd, W, H)
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


def cpu_sample_tile(W, H):
    """Bounded CPU sample of the workload: the central band of rows (1/8 of the frame)."""
    h = max(H // 8, 1)
    return (0, (H - h) // 2, W, h)


def run_cpu(sc, cd, shade, W, H, steps, warmup):
    """Oracle (CPU restatement, OpenMP) on the bounded sample.  Returns (Mrays/s, seconds/step, rays, cores, tile)."""
    from oracle import pyoracle as O
    osc = O.Scene(sc["domains"])
    cam = O.camera_from_desc(cd)
    sh, keep = O.make_shade(lights=sc["lights"], **shade)
    tile = cpu_sample_tile(W, H)
    times, rays = [], 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        fb, st = osc.render(cam, sh, [tile])
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            rays = st.radiance_rays + st.shadow_rays
    sec = float(np.mean(times))
    return rays / sec / 1e6, sec, int(rays), O.lib().orc_num_threads(), tile


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("SPRAY_WORKLOAD", "c2"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--partition", default=os.environ.get("SPRAY_PARTITION", "image"), choices=["image", "insitu", "both"],
                    help="N > 1: image = the reference's --partition image (spray_ooc: all domains on every rank, "
                         "row bands, NCCL gather); insitu = --partition insitu (domains sharded, NCCL min-all-reduce "
                         "on the hit keys); both = headline image, insitu reported beside it")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    sc, cd, shade, label, W, H = workload(args.workload)

    if args.impl == "reference":
        # torchrun exports OMP_NUM_THREADS=1; the CPU arm uses every host core (set before libgomp loads)
        os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
        # The reference's own CPU implementation cannot be built (Embree 2.x, MPI, IceT, OpenGL absent);
        # this arm times the oracle port on all host threads.  Rank 0 only.
        if rank != 0:
            return
        v, sec, rays, cores, tile = run_cpu(sc, cd, shade, W, H, max(args.steps, 1), min(args.warmup, 1))
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic camera rays on the reference's example scene",
                "config": {"workload": label},
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                 "sample": "rows %d..%d of the %dx%d frame (%d rays/step); CPU restatement, NOT Embree" % (tile[1], tile[1] + tile[3], W, H, rays)},
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
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **shade)
    tiles = scenes.image_schedule_tiles(W, H, shade["pixel_samples"], 1, 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    if world > 1:
        from spray_b200 import multigpu
        comm = multigpu.bootstrap_comm(dev, rank, world)
        mode = "image" if args.partition == "both" else args.partition
        if args.workload == "c5" and "SPRAY_PARTITION" not in os.environ and "--partition" not in sys.argv:
            mode = "insitu"                               # C5 is the reference's in-situ configuration
        Renderer = multigpu.ImagePartitionRenderer if mode == "image" else multigpu.InsituRenderer
        mg = Renderer(dev, sc, cam, sh, W, H, rank, world, comm)
        step_dev = mg.frame_device
        step_e2e = mg.frame_host
        st_mg = mg.count_rays()
        rr = torch.tensor([st_mg["radiance_rays"] + st_mg["shadow_rays"]], dtype=torch.int64, device="cuda")
        if mode == "image":
            dist.all_reduce(rr)          # each rank counted its own band
        rays_total = int(rr.item())

        def rays_of():
            return rays_total
    elif args.workload == "c4s":
        # out-of-core: meshes wait in PINNED host memory, the LRU uploads + commits on every miss
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

        def rays_of():
            return ooc_stats["radiance_rays"] + ooc_stats["shadow_rays"]
    else:
        t_build0 = time.perf_counter()
        scenes_by_id = dev.load_scene(sc)
        dev.synchronize()
        build_s = time.perf_counter() - t_build0
        infos = [s_.info() for s_ in scenes_by_id.values()]
        build_dev_ms = float(sum(i.build_ms for i in infos))          # CUDA-event time of every spraygpu_scene_commit
        build_tris = int(sum(i.num_faces for i in infos))
        # steady state of a cache block that is re-loaded (out-of-core LRU miss): same buffers, no allocation
        for s_ in scenes_by_id.values():
            s_.commit()
        rebuild_dev_ms = float(sum(s_.info().build_ms for s_ in scenes_by_id.values()))
        fb_d = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
        fb_h = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory()
        fb_h_np = fb_h.numpy()
        stats_tot = {}

        kernel_events = []                                # (start, end) around the render launch alone, per timed frame

        def step_dev():
            fb_d.zero_()                                  # image_.clear(), spray_renderer.inl:363
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            dev.trace_frame(cam, sh, tiles, fb_d)         # the frame's blocking tiles in one persistent launch
            b.record(stream)
            kernel_events.append((a, b))

        def step_e2e():
            # the public C-ABI call with a pinned HOST frame buffer: camera/shade/tile descriptors go H2D as
            # kernel arguments, every tile's rows come back D2H (copy of tile k overlaps the kernel of tile k+1);
            # the frame is complete after the synchronize
            for t in tiles:
                dev.trace_tile_async(cam, sh, t, fb_h_np)
            dev.synchronize()

        # ray counts (deterministic): one counted frame, also gives N_node / N_tri for the roofline
        fb_d.zero_()
        for t in tiles:
            st = dev.trace_tile(cam, sh, t, fb_d, stats=True, counted=True)
            for k, v in st.as_dict().items():
                stats_tot[k] = stats_tot.get(k, 0) + v

        def rays_of():
            return stats_tot["radiance_rays"] + stats_tot["shadow_rays"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step, steps, warmup):
        for _ in range(warmup):
            step()
        barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        l0 = dev.launch_count()
        for a, b in evs:
            flush.fill_(1)                                # L2 flush between timed iterations (outside the events)
            a.record(stream)
            step()
            b.record(stream)
        barrier()
        launches = dev.launch_count() - l0
        ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_dev, launches = timed(step_dev, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e, _ = timed(step_e2e, max(args.steps // 2, 3), 2)

    rays = rays_of()
    value = rays / (ms_dev * 1e-3) / 1e6
    e2e_v = rays / (ms_e2e * 1e-3) / 1e6

    par_label = "1 GPU, all domains resident"
    insitu_extra = None
    if world > 1:
        par_label = ("image partition x%d (reference --partition image: all domains on every rank, row bands, NCCL gather to rank 0)" % world
                     if mode == "image" else
                     "in-situ domain partition x%d (Morton), NCCL min-all-reduce on 64-bit hit keys + max-all-reduce on occlusion, NCCL image reduce" % world)
        if args.partition == "both":
            dev2 = capi.Device(local_rank, stream=stream.cuda_stream)
            comm2 = multigpu.bootstrap_comm(dev2, rank, world)
            mg2 = multigpu.InsituRenderer(dev2, sc, cam, sh, W, H, rank, world, comm2)
            st2 = mg2.count_rays()
            ms2, _ = timed(mg2.frame_device, max(args.steps // 2, 3), 3)
            r2 = st2["radiance_rays"] + st2["shadow_rays"]
            insitu_extra = {"value": r2 / (ms2 * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms2,
                            "parallelism": "in-situ domain partition x%d, NCCL min-all-reduce on hit keys" % world}
            comm2.close(); dev2.close()
    if rank == 0:
        peak, peak_src = measured_peak_hbm()
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic camera rays on the reference's example scene (random-free)",
                "config": {"workload": label, "tiles_per_frame": len(tiles), "l2_flush_between_steps": True,
                           "rays_per_frame": int(rays), "parallelism": par_label},
                "clocks": clocks, "gpu_launches": int(launches),
                "e2e": {"value": e2e_v, "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": int(len(tiles) * (capi.C.sizeof(capi.Camera) + capi.C.sizeof(capi.Shade))),
                        "d2h_bytes_per_step": int(W * H * 16)}}
        if insitu_extra:
            line["insitu"] = insitu_extra
        if world == 1 and args.workload == "c4s":
            line["config"]["parallelism"] = "1 GPU, out-of-core: %d cache blocks for %d domains" % (cap, len(sc["domains"]))
            mesh_bytes = int(sum(d["verts"].nbytes + d["faces"].nbytes + d["rgb"].nbytes for d in sc["domains"]) // len(sc["domains"]))
            line["e2e"]["h2d_bytes_per_step"] += int(misses_per_frame) * mesh_bytes     # meshes re-uploaded by the LRU
            line["ooc"] = {"cache_blocks": cap, "mesh_bytes_per_domain": mesh_bytes, "domains": len(sc["domains"]), "misses_first_frame": int(misses_per_frame),
                           "misses_total": int(cache.misses), "hits_total": int(cache.hits),
                           "note": "every miss = H2D of the domain's mesh from pinned memory + spraygpu_scene_commit"}
        elif world == 1:
            nn, nt, nh = stats_tot["nodes_visited"], stats_tot["tris_tested"], stats_tot["hits"]
            nrays = rays
            # algorithmic bytes: per ray 16 B frame-buffer + 80 B/node + 48 B/triangle (+48 B attribute record per hit)
            bytes_frame = 16 * W * H + 80 * nn + 48 * nt + 48 * nh
            nl = max(int(launches) // max(args.steps, 1), 1)        # render launches per frame (1: spraygpu_trace_frame)
            per_launch = bytes_frame / nl
            # the kernel's own duration: events around the launch inside every TIMED frame (the step also clears the image)
            kern_ms = float(np.mean([a.elapsed_time(b) for a, b in kernel_events[-args.steps:]]))
            dur_s = kern_ms * 1e-3 / nl
            achieved = per_launch / dur_s / 1e9
            line["roofline"] = {"bound": "hbm", "kernel": "render_tile_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                                "frac": achieved / peak, "traffic": NCU_TRAFFIC.get(args.workload), "peak_source": peak_src,
                                "traffic_source": NCU_TRAFFIC_SOURCE if args.workload in NCU_TRAFFIC else None,
                                "algorithmic_bytes_per_launch": per_launch, "launch_ms": dur_s * 1e3,
                                "nodes_per_ray": nn / nrays, "tris_per_ray": nt / nrays, "bytes_per_ray": bytes_frame / nrays,
                                "domain_tests_per_ray": stats_tot["domain_tests"] / nrays}
            line["config"]["build_s_all_domains"] = build_s
            # per-domain acceleration-structure build (= rtcCommit): device time of the commits alone; the wall-clock
            # figure above also holds the H2D upload of the meshes from pageable host memory
            line["build"] = {"triangles": build_tris, "first_commit_device_ms": build_dev_ms,
                             "recommit_device_ms": rebuild_dev_ms,
                             "mtris_per_s": build_tris / max(rebuild_dev_ms, 1e-9) / 1e3,
                             "algorithmic_gb_per_s": build_tris * 300.0 / max(rebuild_dev_ms, 1e-9) / 1e6,
                             "note": "first commit includes cudaMalloc of the domain's buffers; mtris_per_s is the re-commit"}
            if not args.no_cpu_baseline:
                v, sec, crays, cores, tile = run_cpu(sc, cd, shade, W, H, 3, 1)
                line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                        "sample": "rows %d..%d of the frame (%d rays), 3 steps; CPU restatement (oracle O2), NOT Embree" % (tile[1], tile[1] + tile[3], crays)}
        print(json.dumps(line))
    if world > 1:
        comm.close()
        dist.destroy_process_group()
    dev.close()


if __name__ == "__main__":
    main()
