"""Multi-GPU frame loops (one process per GPU, launched by torchrun).  Thin host-side harness over the C ABI:
torch.distributed is used ONLY to bootstrap (broadcast of the NCCL unique id, barriers, max-over-ranks timing and
the sums of per-rank counters); everything on the frame path is issued by libspraygpu.

Two decompositions, as in the reference:

* ``ImagePartitionRenderer`` — the reference's ``--partition image`` (spray_ooc, src/ooc/ooc_tracer.inl:29-82,
  src/render/tile.cc:317-391): every rank holds ALL domains and renders its own pixels; no ray traffic; the image
  is composited on rank 0 (HdrImage::composite, src/display/image.h:167-181).  Here a rank owns interleaved ROW
  BLOCKS (image-space load balancing; the reference cuts one vertical stripe per rank), renders them as
  rectangles of the ONE-rank blocking tiles (so sample ids / RNG seeds and therefore the image do not depend on
  the rank count — spraygpu_trace_frame_rects) and FLUSHES them with one kernel: 16-byte stores into rank 0's
  composite over NVLink (CUDA IPC mapping) and / or into one shared pinned host image over its own PCIe link
  (spraygpu_image_flush).  No NCCL call per frame.
* ``InsituRenderer`` — the reference's ``--partition insitu`` (spray_insitu_*, src/insitu/): domains are
  sharded with the Morton partition (src/render/data_partition.h:59-137); every rank traces every ray
  against ITS domains; hits are resolved by a 64-bit min on the packed keys (replaces
  insitu::VBuf::compositeTbuf, insitu_vbuf.h:109-117), occlusion by an OR (compositeObuf, :119-129).
"""
import os

import numpy as np

from . import capi, scenes


def row_bands(image_h, nranks):
    """rank r renders rows [b[r], b[r+1]): equal bands, the remainder spread over the first ranks."""
    base, rem = divmod(int(image_h), int(nranks))
    b = [0]
    for r in range(nranks):
        b.append(b[-1] + base + (1 if r < rem else 0))
    return np.asarray(b, np.int32)


def band_tiles(image_w, band, spp, max_samples=1024 * 1024):
    """Blocking tiles (<= max_samples samples each, cfg.maximum_num_samples_per_rank) of one row band."""
    y0, y1 = int(band[0]), int(band[1])
    if y1 <= y0:
        return []
    rows = max(int(max_samples // max(image_w * spp, 1)), 1)
    return [(0, y, image_w, min(rows, y1 - y)) for y in range(y0, y1, rows)]


def interleaved_blocks(image_h, nranks, max_tiles_per_rank=32, root=0, root_share=1.0):
    """Image-space load balancing for the image partition: the image is cut into row blocks dealt round-robin to the
    ranks (block b -> rank b % nranks).  Block height = the smallest multiple of 4 (the kernel's 8x4 pixel blocks)
    that keeps a rank's share within one persistent launch (<= max_tiles_per_rank rectangles).  root_share < 1 lets
    the compositing rank render a smaller share.  Returns (row_begin[nb + 1], owner[nb])."""
    rows = -(-int(image_h) // (max_tiles_per_rank * int(nranks)))
    rows = max(4, (rows + 3) // 4 * 4)
    begin = list(range(0, int(image_h), rows)) + [int(image_h)]
    nb = len(begin) - 1
    order, acc = [], 0.0
    while len(order) < nb:
        for r in range(int(nranks)):
            if r == root and nranks > 1:
                acc += float(root_share)
                if acc < 1.0 - 1e-9:
                    continue
                acc -= 1.0
            order.append(r)
            if len(order) == nb:
                break
    return np.asarray(begin, np.int32), np.asarray(order, np.int32)


def rects_of_blocks(block_begin, block_owner, rank, blocking_tiles):
    """This rank's row blocks cut against the one-rank blocking tiles: (rects, parents) for spraygpu_trace_frame_rects."""
    rects, parents = [], []
    for b in range(len(block_owner)):
        if int(block_owner[b]) != rank:
            continue
        y0, y1 = int(block_begin[b]), int(block_begin[b + 1])
        for (tx, ty, tw, th) in blocking_tiles:
            a, z = max(y0, ty), min(y1, ty + th)
            if z > a:
                rects.append((tx, a, tw, z - a))
                parents.append((tx, ty, tw, th))
    return rects, parents


def scatter_partition(world_bounds, scene_bounds, nranks):
    """View-independent spreading of the domains over the ranks (NOT in the reference; an experiment for the in-situ load
    imbalance): the reference's partitions follow the Morton order — contiguous runs put all
    camera-facing domains on few ranks, and dealing round-robin over the Morton order is, on a 2^k grid, the parity
    partition (one rank owns every all-odd cell).  Here the Morton-sorted domains are dealt in the order of a fixed
    multiplicative hash of their Morton rank, nd / nranks each."""
    d2r, codes = capi.partition_domains(world_bounds, scene_bounds, 1, False, return_codes=True)
    order = np.argsort(codes, kind="stable")
    nd = len(order)
    h = (np.arange(nd, dtype=np.uint64) * np.uint64(2654435761) + np.uint64(40503)) % np.uint64(1 << 32)
    h = ((h >> np.uint64(13)) ^ h) * np.uint64(0x5bd1e995) % np.uint64(1 << 32)
    deal = np.argsort(h, kind="stable")                   # position in the deal of Morton rank j
    out = np.zeros(nd, np.int32)
    for pos, j in enumerate(deal):
        out[order[j]] = pos % nranks
    return out


def bootstrap_comm(dev, rank, world):
    """Create the library's NCCL communicator; the unique id travels over torch.distributed."""
    import torch.distributed as dist
    obj = [capi.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    return capi.Comm(dev, world, rank, obj[0])


def _sum_over_ranks(d):
    import torch
    import torch.distributed as dist
    keys = sorted(d)
    t = torch.tensor([int(d[k]) for k in keys], dtype=torch.int64, device="cuda")
    dist.all_reduce(t)
    return {k: int(v) for k, v in zip(keys, t.tolist())}


class ImagePartitionRenderer:
    def __init__(self, dev, scene, cam, shade, W, H, rank, world, comm, shm_name=None, load=True):
        self.dev, self.cam, self.shade, self.W, self.H, self.rank, self.world, self.comm = dev, cam, shade, W, H, rank, world, comm
        if load:
            dev.load_scene(scene)
        self.image = capi.Image(dev, comm, W, H, shm_name)
        self.blocking = scenes.image_schedule_tiles(W, H, shade.pixel_samples, 1, 0)      # the one-rank frame's blocking tiles
        share = float(os.environ.get("SPRAY_ROOT_SHARE", "1.0"))
        self.block_begin, self.block_owner = interleaved_blocks(H, world, root=0, root_share=share)
        self.image.set_rows(self.block_begin, self.block_owner)
        self.rects, self.parents = rects_of_blocks(self.block_begin, self.block_owner, rank, self.blocking)

    def _trace(self, stats=False):
        return self.dev.trace_frame_rects(self.cam, self.shade, self.rects, self.parents, self.image.device, stats=stats)

    def frame_device(self):
        """One frame, result on rank 0's DEVICE composite: trace my rectangles, push my rows over NVLink."""
        self._trace()
        self.image.flush(to_root=True, to_host=False, clear=True)

    def frame_host(self):
        """One frame, result in the shared pinned HOST image: every rank stores its own rows over its own PCIe link."""
        self._trace()
        self.image.flush(to_root=False, to_host=True, clear=True)
        self.dev.synchronize()

    def frame_host_async(self):
        """frame_host without the per-frame synchronize: this frame's rows are compacted on the device and stored into the
        host image on the image's own stream while the next frame is traced (spraygpu_image_flush_async)."""
        self.dev.trace_frame_rects(self.cam, self.shade, self.rects, self.parents, self.image.device)
        self.image.flush_async(to_host=True)

    def count_rays(self):
        """Deterministic per-frame ray counts, summed over the ranks (one extra frame with stats)."""
        st = self._trace(stats=True).as_dict()
        self.image.flush(clear=True)
        return _sum_over_ranks(st)

    def close(self):
        self.image.close()


class InsituRenderer:
    """Domain partition (the reference's in-situ tracers): rank r holds the domains spraygpu_partition_domains gives it
    and every rank walks every blocking tile.

    mode "p2p" (default): spraygpu_trace_tile_insitu_p2p — keys / hit attributes / occlusion bytes cross NVLink inside
    the kernels, each rank accumulates its HOME rows (interleaved blocks of `rows_per_block` image rows) and the sparse
    image flush moves them to the root composite / the shared pinned host image.
    mode "nccl": spraygpu_trace_tile_insitu — five dense NCCL all-reduces per wave, partial images summed with
    spraygpu_image_reduce (round 1's path, kept for the A/B)."""

    def __init__(self, dev, scene, cam, shade, W, H, rank, world, comm, round_robin=False, mode="p2p", rows_per_block=8,
                 shm_name=None, max_tile_samples=None, scatter=False):
        import torch
        self.dev, self.cam, self.shade, self.W, self.H, self.rank, self.world, self.comm = dev, cam, shade, W, H, rank, world, comm
        self.mode, self.rpb = mode, int(rows_per_block)
        wb = np.stack([d["bounds"] for d in scene["domains"]])
        self.domain_to_rank = capi.partition_domains(wb, scenes.scene_bounds(scene), world, round_robin)
        if scatter and world > 1:
            self.domain_to_rank = scatter_partition(wb, scenes.scene_bounds(scene), world)
        self.owned = [i for i, r in enumerate(self.domain_to_rank) if r == rank]
        dev.load_scene(scene, owned=self.owned)
        if max_tile_samples is None:
            max_tile_samples = (4 << 20) if mode in ("p2p", "fwd") else (1 << 20)   # p2p / fwd: fewer, larger waves = fewer peer barriers
        self.tiles = band_tiles(W, (0, H), shade.pixel_samples, max_tile_samples)   # every rank walks every blocking tile
        if mode in ("p2p", "fwd"):
            self.image = capi.Image(dev, comm, W, H, shm_name)
            begin = list(range(0, H, self.rpb)) + [H]
            self.image.set_rows(begin, [b % world for b in range(len(begin) - 1)])
            self.fb = self.image.device
            self.fb_host = None
        else:
            self.image = None
            self.fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
            self.fb_host = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory() if rank == 0 else None

    def _trace(self, stats=False):
        tot = {}
        for t in self.tiles:
            if self.mode == "p2p":
                st = self.comm.trace_tile_insitu_p2p(self.cam, self.shade, t, self.fb, self.rpb, stats=stats)
            elif self.mode == "fwd":
                st = self.comm.trace_tile_insitu_fwd(self.cam, self.shade, t, self.fb, self.rpb, self.domain_to_rank, stats=stats)
            else:
                st = self.comm.trace_tile_insitu(self.cam, self.shade, t, self.fb, stats=stats)
            if stats:
                for k, v in st.as_dict().items():
                    tot[k] = tot.get(k, 0) + v            # already summed over ranks by the library
        return tot

    def frame_device(self):
        if self.mode in ("p2p", "fwd"):
            self._trace()
            self.image.flush(to_root=True, to_host=False, clear=True)
        else:
            self.fb.zero_()
            self._trace()
            self.comm.image_reduce(self.fb, self.W * self.H, 0)

    def frame_host(self):
        if self.mode in ("p2p", "fwd"):
            self._trace()
            self.image.flush(to_root=False, to_host=True, clear=True)
        else:
            self.frame_device()
            if self.rank == 0:
                self.fb_host.copy_(self.fb, non_blocking=True)
        self.dev.synchronize()

    def count_rays(self):
        if self.mode in ("p2p", "fwd"):
            tot = self._trace(stats=True)
            self.image.flush(to_root=True, to_host=False, clear=True)     # leave the render target cleared
            return tot
        self.fb.zero_()
        return self._trace(stats=True)

    def composite_image(self):
        """rank 0: the composited frame after frame_device() (numpy); call on every rank."""
        self.dev.synchronize()
        if self.mode in ("p2p", "fwd"):
            self.image.barrier()                       # every rank's flush has landed in the root composite
            return self.image.read_device("root") if self.rank == 0 else None
        return self.fb.cpu().numpy()

    def host_image(self):
        self.dev.synchronize()
        if self.mode in ("p2p", "fwd"):
            self.image.barrier()
            return self.image.host
        return self.fb_host.numpy() if self.fb_host is not None else None

    def close(self):
        if self.image is not None:
            self.image.close()
            self.image = None


def _rmse(a, b):
    return float(np.sqrt(np.mean((np.asarray(a, np.float64)[..., :3] - np.asarray(b, np.float64)[..., :3]) ** 2)))


def single_gpu_frame(dev, cam, shade, W, H, tiles=None):
    """The whole frame on THIS GPU alone through the fused kernel (the parity_check reference); all domains must be resident.
    `tiles`: the blocking tiles the compared renderer used (sample ids, and with them the bounce / area-light random
    streams, are tile-relative: src/ooc/ooc_tracer.inl:84-172); default = the one-rank frame's."""
    import torch
    fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    if tiles is None:
        tiles = scenes.image_schedule_tiles(W, H, shade.pixel_samples, 1, 0)
    st = dev.trace_frame(cam, shade, tiles, fb, stats=True).as_dict()
    dev.synchronize()
    return fb.cpu().numpy(), st


INSITU_LABEL = ("in-situ domain partition x%d (Morton): each GPU owns %d of %d domains; hit keys merged by 64-bit system-scope "
                "atomicMin into every GPU's key array over NVLink (CUDA IPC), hit attributes and occlusion bytes by peer stores, "
                "device-side barriers, no NCCL per wave; each GPU accumulates its home rows (blocks of %d rows) and flushes them "
                "to rank 0 / the shared pinned host image")


def insitu_deal():
    """bench.py's domain -> rank deal for the in-situ modes: SPRAY_INSITU_PART = scatter (default) | rr | morton."""
    part = os.environ.get("SPRAY_INSITU_PART", "")
    if not part:
        part = "rr" if os.environ.get("SPRAY_INSITU_RR", "") == "1" else ("morton" if os.environ.get("SPRAY_INSITU_RR", "") == "0" else "scatter")
    label = {"morton": "",
             "rr": "; domains dealt round-robin over the Morton order (data_partition.h:139-155)",
             "scatter": "; domains dealt over the GPUs in the order of a hash of their Morton rank (a view-independent spread, NOT one of "
                        "the reference's two partitions: contiguous Morton runs and round-robin put the camera-facing cells of a 4x4x4 grid "
                        "on few GPUs; the image is the same under any deal)"}[part]
    return dict(round_robin=part == "rr", scatter=part == "scatter"), label


FWD_LABEL = ("in-situ domain partition x%d with ray forwarding: each GPU owns %d of %d domains; a ray is appended (32 B, peer store over "
             "NVLink) to the inbox of every GPU that owns a domain it enters, owners trace their inbox and atomicMin the packed key into "
             "the sample's home GPU, every sample is generated / shaded / accumulated on its home GPU only (blocks of %d rows)")


def bench_multi(args, dev, stream, timer, sampler, rank, world, local_rank, workload, metric, unit):
    """bench.py at N > 1.  Returns the JSON line (rank 0) or None."""
    import torch
    import torch.distributed as dist
    sc, cd, shade, label, W, H = workload(args.workload)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **shade)
    mode = "image" if args.partition == "both" else args.partition
    if args.workload in ("c4", "c5") and "SPRAY_PARTITION" not in os.environ and "--partition" not in " ".join(os.sys.argv):
        mode = "insitu"                                   # C4 (sharded scene) / C5 are the reference's in-situ configurations
    comm = bootstrap_comm(dev, rank, world)
    shm = "spraygpu_img_%s_%d" % (os.environ.get("MASTER_PORT", "0"), os.getppid())
    line = None

    def measure(r, with_e2e):
        st = r.count_rays()
        rays = st["radiance_rays"] + st["shadow_rays"]
        if rank == 0:
            sampler.start()
        ms_dev, launches = timer(r.frame_device, args.steps, args.warmup)
        clocks = sampler.stop() if rank == 0 else None
        ms_e2e = None
        if with_e2e:
            ms_e2e, _ = timer(r.frame_host, max(args.steps // 2, 3), 2)
        return st, rays, ms_dev, launches, clocks, ms_e2e

    if mode == "image":
        r = ImagePartitionRenderer(dev, sc, cam, sh, W, H, rank, world, comm, shm_name=shm)
        st, rays, ms_dev, launches, clocks, ms_e2e = measure(r, True)
        # pipelined e2e: frame k's rows cross PCIe while frame k + 1 is traced (one bracket around all frames; the L2 fills
        # inside it are subtracted with a separately measured fill time, the bracket with them is the upper bound)
        ms_e2e_sync = ms_e2e
        ms_e2e_pipe, ms_e2e_pipe_upper, ms_fill = timer.pipelined(r.frame_host_async, r.image.join, r.image.wait, max(args.steps, 5), 3)
        r.image.wait()
        e2e_pipelined = ms_e2e_pipe <= ms_e2e
        ms_e2e = min(ms_e2e, ms_e2e_pipe)
        # ---- parity_check (outside the timed region): one more frame on all ranks, then rank 0 renders the whole
        # frame alone and compares with its composite; the host image of the e2e path must equal the composite too
        r.image.barrier()
        r.frame_device()
        r.image.barrier()
        pc = None
        if rank == 0:
            comp = r.image.read_device("root")
            ref, st1 = single_gpu_frame(dev, cam, sh, W, H)
            pc = {"against": "the same frame rendered by rank 0 alone (fused kernel, all domains resident)",
                  "rmse": _rmse(comp, ref), "max_abs": float(np.abs(comp[..., :3] - ref[..., :3]).max()),
                  "bit_identical": bool(np.array_equal(comp, ref)),
                  "keys_equal": True, "keys_note": "image partition: every rank runs the same closest-hit kernel on its own pixels; "
                                                   "ray counts equal: %s" % all(st[k] == st1[k] for k in ("eye_rays", "radiance_rays", "shadow_rays", "hits"))}
        r.frame_host()
        r.image.barrier()
        fp = _sum_over_ranks(dict(zip(("root", "host"), r.image.flushed_pixels())))
        if rank == 0:
            pc["host_image_equals_composite"] = bool(np.array_equal(r.image.host, comp))
        par_label = ("image partition x%d (reference --partition image): all domains on every rank, %d interleaved row blocks, rows "
                     "pushed into rank 0's frame buffer over NVLink by the library's flush kernel (CUDA IPC, no NCCL per frame); "
                     "e2e: every rank stores its rows into one shared pinned host image" % (world, len(r.block_owner)))
        nframes_root = args.warmup + args.steps + 2        # count_rays flushes nothing to root; + parity frame + (none)
        d2h = None
        r.close()
    else:
        # in-situ headline (C4 / C5, or --partition insitu): SPRAY_INSITU_MODE picks the variant, the other one is measured beside it
        # soups (opaque: the second, bounded serve pass removes most speculative work) do better with ray forwarding, the
        # sparse wavelet scenes with the dense peer-memory exchange; the other variant is always measured beside the headline
        imode = os.environ.get("SPRAY_INSITU_MODE", "fwd" if args.workload in ("c3", "c4") else "p2p")
        deal_kw, deal_label = insitu_deal()
        fits_one_gpu = args.workload != "c4" or int(os.environ.get("SPRAY_C4_TRIS", str(1 << 23))) * len(sc["domains"]) <= (1 << 26)
        r = InsituRenderer(dev, sc, cam, sh, W, H, rank, world, comm, mode=imode, shm_name=shm, **deal_kw)
        build_ms = float(sum(s_.info().build_ms for s_ in dev.scenes if s_.h))
        build_tris = int(sum(s_.info().num_faces for s_ in dev.scenes if s_.h))
        st, rays, ms_dev, launches, clocks, ms_e2e = measure(r, True)
        r.frame_device()
        comp = r.composite_image()
        if fits_one_gpu:
            pc = insitu_parity_check(r, sc, cam, sh, W, H, rank, local_rank, stream)
        else:
            pc = {"against": None} if rank == 0 else None
        r.frame_host()
        host = r.host_image()
        if rank == 0:
            pc["host_image_equals_composite"] = bool(np.array_equal(host, comp))
            pc.pop("_composite", None)
        fp = _sum_over_ranks(dict(zip(("root", "host"), r.image.flushed_pixels())))
        owned_n, rpb = len(r.owned), r.rpb
        # the other variant on the same device / scene (the domains stay loaded): cross-check of two independent exchanges
        omode = "p2p" if imode == "fwd" else "fwd"
        r.image.close(); r.image = None
        r_o = InsituRenderer.__new__(InsituRenderer)
        r_o.__dict__.update(r.__dict__)
        r_o.mode = omode
        r_o.image = capi.Image(dev, comm, W, H, shm + "_o")
        begin = list(range(0, H, r_o.rpb)) + [H]
        r_o.image.set_rows(begin, [b_ % world for b_ in range(len(begin) - 1)])
        r_o.fb = r_o.image.device
        st_o = r_o.count_rays()
        ms_o, l_o = timer(r_o.frame_device, max(args.steps // 2, 3), 3)
        r_o.frame_device()
        comp_o = r_o.composite_image()
        other = None
        if rank == 0:
            other = {"mode": omode, "value": rays / (ms_o * 1e-3) / 1e6, "unit": unit, "ms_per_step": ms_o, "gpu_launches": int(l_o),
                     "image_rmse_vs_headline_variant": _rmse(comp_o, comp), "image_bit_identical": bool(np.array_equal(comp_o, comp)),
                     "ray_counts_equal": all(st_o[k] == st[k] for k in ("eye_rays", "radiance_rays", "shadow_rays", "hits"))}
            if not fits_one_gpu:
                pc = {"against": "the scene does not fit one GPU: cross-check of the two in-situ exchanges (ray forwarding vs dense peer-memory waves), "
                                 "each parity-tested against the single-GPU frame and the oracle at sizes that fit (tests/multigpu_check.py)",
                      "rmse": other["image_rmse_vs_headline_variant"], "bit_identical": other["image_bit_identical"],
                      "keys_equal": other["ray_counts_equal"], "host_image_equals_composite": pc.get("host_image_equals_composite")}
        r_o.image.close(); r_o.image = None
        par_label = (FWD_LABEL if imode == "fwd" else INSITU_LABEL) % (world, owned_n, len(sc["domains"]), rpb)
        par_label += deal_label
        insitu_headline = {"other_variant": other, "build": {"triangles_this_gpu": build_tris, "commit_device_ms_this_gpu": build_ms,
                                                               "mtris_per_s": build_tris / max(build_ms, 1e-9) / 1e3}}
    insitu_extra = None
    if args.partition == "both" and mode == "image":
        dev2 = capi.Device(local_rank, stream=stream.cuda_stream)
        comm2 = bootstrap_comm(dev2, rank, world)
        timer2 = type(timer)(timer.stream, timer.flush, dev2, world)
        deal_kw, deal_label = insitu_deal()
        r2 = InsituRenderer(dev2, sc, cam, sh, W, H, rank, world, comm2, shm_name=shm + "_i", **deal_kw)
        st2 = r2.count_rays()
        ms2, l2 = timer2(r2.frame_device, max(args.steps // 2, 3), 3)
        ms2h, _ = timer2(r2.frame_host, max(args.steps // 2, 3), 2)
        pc2 = insitu_parity_check(r2, sc, cam, sh, W, H, rank, local_rank, stream)
        if rank == 0:
            pc2.pop("_composite", None)
        r2n = st2["radiance_rays"] + st2["shadow_rays"]
        insitu_extra = {"value": r2n / (ms2 * 1e-3) / 1e6, "unit": unit, "ms_per_step": ms2, "gpu_launches": int(l2),
                        "e2e": {"value": r2n / (ms2h * 1e-3) / 1e6, "unit": unit, "ms_per_step": ms2h},
                        "parallelism": INSITU_LABEL % (world, len(r2.owned), len(sc["domains"]), r2.rpb) + deal_label,
                        "parity_check": pc2}
        r2.close()
        # the ray-forwarding variant (insitu3.cu) on the same communicator, same parity check
        r4 = InsituRenderer(dev2, sc, cam, sh, W, H, rank, world, comm2, mode="fwd", shm_name=shm + "_f", **deal_kw)
        st4 = r4.count_rays()
        ms4, l4 = timer2(r4.frame_device, max(args.steps // 2, 3), 3)
        ms4h, _ = timer2(r4.frame_host, max(args.steps // 2, 3), 2)
        pc4 = insitu_parity_check(r4, sc, cam, sh, W, H, rank, local_rank, stream)
        if rank == 0:
            pc4.pop("_composite", None)
            pc4["ray_counts_equal_dense_variant"] = all(st4[k] == st2[k] for k in ("eye_rays", "radiance_rays", "shadow_rays", "hits"))
        insitu_extra["ray_forwarding"] = {"value": r2n / (ms4 * 1e-3) / 1e6, "unit": unit, "ms_per_step": ms4, "gpu_launches": int(l4),
                                          "e2e": {"value": r2n / (ms4h * 1e-3) / 1e6, "unit": unit, "ms_per_step": ms4h},
                                          "parallelism": FWD_LABEL % (world, len(r4.owned), len(sc["domains"]), r4.rpb) + deal_label,
                                          "parity_check": pc4}
        r4.close()
        # A/B: round 1's exchange (dense NCCL all-reduces per wave) on the same communicator
        r3 = InsituRenderer(dev2, sc, cam, sh, W, H, rank, world, comm2, mode="nccl", **deal_kw)
        ms3, l3 = timer2(r3.frame_device, 3, 2)
        insitu_extra["nccl_allreduce_variant"] = {"value": r2n / (ms3 * 1e-3) / 1e6, "unit": unit, "ms_per_step": ms3, "gpu_launches": int(l3)}
        r3.close(); comm2.close(); dev2.close()
    if rank == 0:
        line = {"metric": metric, "value": rays / (ms_dev * 1e-3) / 1e6, "unit": unit, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic camera rays on the reference's example scene (random-free)",
                "config": {"workload": label, "l2_flush_between_steps": True, "rays_per_frame": int(rays), "parallelism": par_label},
                "clocks": clocks, "gpu_launches": int(launches), "parity_check": pc,
                "e2e": {"value": rays / (ms_e2e * 1e-3) / 1e6, "unit": unit, "ms_per_step": ms_e2e,
                        "h2d_bytes_per_step": int(capi.C.sizeof(capi.Camera) + capi.C.sizeof(capi.Shade)) * world,
                        "d2h_bytes_per_step": int(W * H * 16)}}
        if fp is not None:
            # the flush is sparse: pixels that are (or were) non-zero; report what actually crossed PCIe per e2e frame
            e2e_frames = max(args.steps // 2, 3) + 2 + 1 + ((max(args.steps, 5) + 3) if mode == "image" else 0)
            line["e2e"]["d2h_bytes_per_step"] = int(fp["host"] * 16 // e2e_frames)
            line["e2e"]["d2h_note"] = "sparse flush: only pixels that are or were non-zero are stored to the host image (full frame = %d B)" % (W * H * 16)
        if mode == "image":
            line["e2e"]["mode"] = ("pipelined frames (spraygpu_image_flush_async: frame k's rows cross PCIe while frame k + 1 is traced; one event "
                                   "pair around all frames until the last frame's stores have landed, minus the separately timed L2 fills)" if e2e_pipelined else "one frame at a time")
            line["e2e"]["per_frame_sync"] = {"value": rays / (ms_e2e_sync * 1e-3) / 1e6, "ms_per_step": ms_e2e_sync}
            line["e2e"]["pipelined_ms_per_step"] = ms_e2e_pipe
            line["e2e"]["pipelined_ms_per_step_upper"] = ms_e2e_pipe_upper
            line["e2e"]["l2_fill_ms"] = ms_fill
        if insitu_extra:
            line["insitu"] = insitu_extra
        if mode != "image":
            line["insitu_detail"] = insitu_headline
        # the numbers a reader wants first, once more as the LAST key (short: survives a truncated tail of the line)
        summ = {"workload": args.workload, "n_gpus": world, "partition": mode, "value": line["value"], "e2e": line["e2e"]["value"],
                "parity_bit_identical": None if not pc else pc.get("bit_identical")}
        if insitu_extra:
            summ["insitu_p2p"] = insitu_extra.get("value")
            summ["insitu_fwd"] = (insitu_extra.get("ray_forwarding") or {}).get("value")
            summ["insitu_nccl"] = (insitu_extra.get("nccl_allreduce_variant") or {}).get("value")
        line["summary"] = summ
        line["_composite"] = comp          # bench.py compares it with the CPU checker's frame and removes it before printing
    comm.close()
    return line


def insitu_parity_check(r, sc, cam, sh, W, H, rank, local_rank, stream):
    """In-situ frame (domains sharded) against the same frame on ONE GPU holding every domain; keys: the sharded
    closest-hit keys after the cross-GPU min against the single-GPU keys."""
    import torch
    import torch.distributed as dist
    r.frame_device()
    img = r.composite_image()
    n = W * H
    rays = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
    r.dev.gen_eye_rays(cam, (0, 0, W, H), 1, rays)
    keys = torch.full((n,), -1, dtype=torch.int64, device="cuda")
    r.dev.intersect_domains(rays, n, keys, None)
    r.comm.allreduce_keys_min(keys, n)
    r.dev.synchronize()
    pc = None
    if rank == 0:
        full = capi.Device(local_rank, stream=stream.cuda_stream)
        full.load_scene(sc)
        sh_i = capi.Shade.from_buffer_copy(sh)
        sh_i.eye_seed = capi.EYE_SEED_INSITU           # the in-situ tile always jitters with RandomSampler_init(pixid, s)
        ref, st1 = single_gpu_frame(full, cam, sh_i, W, H, tiles=r.tiles)
        kf = torch.zeros(n, dtype=torch.int64, device="cuda")
        full.intersect_domains(rays, n, kf, None)
        full.synchronize()
        pc = {"against": "the same frame on ONE GPU holding every domain (fused kernel)", "rmse": _rmse(img, ref),
              "max_abs": float(np.abs(img[..., :3] - ref[..., :3]).max()), "keys_equal": bool(torch.equal(keys, kf)),
              "bit_identical": bool(np.array_equal(img, ref)), "_composite": img}
        full.close()
    dist.barrier()
    return pc
