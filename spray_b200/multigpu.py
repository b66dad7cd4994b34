"""Multi-GPU frame loops (one process per GPU, launched by torchrun).  Thin host-side harness over the C ABI:
torch.distributed is used ONLY to bootstrap (broadcast of the NCCL unique id, barriers, max-over-ranks
timing); every data-path collective is issued by libspraygpu itself on its own NCCL communicator.

Two decompositions, as in the reference:

* ``ImagePartitionRenderer`` — the reference's ``--partition image`` (spray_ooc, src/ooc/ooc_tracer.inl:29-82,
  src/render/tile.cc:317-391): every rank holds ALL domains and renders its own pixels; no ray traffic;
  the image is composited on rank 0 (HdrImage::composite, src/display/image.h:167-181).  Ranks own
  contiguous ROW bands (the reference cuts vertical stripes; rows are contiguous in memory, so the
  composite is one NCCL send per rank instead of a whole-image reduce).
* ``InsituRenderer`` — the reference's ``--partition insitu`` (spray_insitu_*, src/insitu/): domains are
  sharded with the Morton partition (src/render/data_partition.h:59-137); every rank traces every ray
  against ITS domains; hits are resolved by an NCCL min-all-reduce on the packed 64-bit keys
  (replaces insitu::VBuf::compositeTbuf, insitu_vbuf.h:109-117), occlusion by a max-all-reduce
  (compositeObuf, :119-129); partial images are summed on rank 0.
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


def interleaved_blocks(image_h, nranks, max_tiles_per_rank=4, root=0, root_share=1.0):
    """Image-space load balancing for the image partition (default from 4 GPUs up).  A fine cut (32 blocks per rank)
    lost badly to contiguous bands (one NCCL send per block); 4 blocks per rank is the measured compromise.
    The image is cut into row blocks dealt round-robin to the ranks (block b -> rank
    b % nranks).  Block height = the smallest multiple of 4 (the kernel's 8x4 pixel blocks) that keeps a rank's
    share within one persistent launch (<= max_tiles_per_rank tiles).  Returns (row_begin[nb + 1], owner[nb])."""
    rows = -(-int(image_h) // (max_tiles_per_rank * int(nranks)))
    rows = max(4, (rows + 3) // 4 * 4)
    begin = list(range(0, int(image_h), rows)) + [int(image_h)]
    # root_share < 1: the compositing rank renders a smaller share, so its receives are posted (and the other ranks'
    # rows start to flow) while the slower ranks still render; 0 makes it a pure compositor
    owner, b, acc = [], 0, 0.0
    nb = len(begin) - 1
    order = []
    cycle = 0
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
        cycle += 1
    owner = order
    return np.asarray(begin, np.int32), np.asarray(owner, np.int32)


def bootstrap_comm(dev, rank, world):
    """Create the library's NCCL communicator; the unique id travels over torch.distributed."""
    import torch.distributed as dist
    obj = [capi.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    return capi.Comm(dev, world, rank, obj[0])


class ImagePartitionRenderer:
    def __init__(self, dev, scene, cam, shade, W, H, rank, world, comm, interleave=None):
        import torch
        self.dev, self.cam, self.shade, self.W, self.H, self.rank, self.world, self.comm = dev, cam, shade, W, H, rank, world, comm
        dev.load_scene(scene)
        if interleave is None:
            # measured (profiles/bench_r01q_*): contiguous bands win at 2 GPUs (C2 10401 vs 9475 Mrays/s), 4 interleaved
            # blocks per rank win at 8 (C2 18343 vs 16645, C3 4279 vs 3654)
            env = os.environ.get("SPRAY_INTERLEAVE")
            interleave = (world >= 4) if env is None else env != "0"
        self.interleave = interleave
        self.bands = row_bands(H, world)
        if interleave:
            share = float(os.environ.get("SPRAY_ROOT_SHARE", "1.0"))
            self.block_begin, self.block_owner = interleaved_blocks(H, world, root=0, root_share=share)
            self.tiles = []
            self.my_blocks = [b for b in range(len(self.block_owner)) if self.block_owner[b] == rank]
            for b in range(len(self.block_owner)):
                if self.block_owner[b] == rank:
                    self.tiles += band_tiles(W, self.block_begin[b:b + 2], shade.pixel_samples)
        else:
            self.tiles = band_tiles(W, self.bands[rank:rank + 2], shade.pixel_samples)
        self.fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
        self.fb_host = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory() if rank == 0 else None

    def frame_device(self):
        if self.interleave:
            for b in self.my_blocks:                             # image_.clear() on the rows this rank renders
                self.fb[int(self.block_begin[b]):int(self.block_begin[b + 1])].zero_()
            for t0 in range(0, len(self.tiles), 32):             # this rank's row blocks, 32 per persistent launch
                self.dev.trace_frame(self.cam, self.shade, self.tiles[t0:t0 + 32], self.fb)
            self.comm.image_gather_blocks(self.fb, self.W, self.block_begin, self.block_owner, 0)
            return
        y0, y1 = int(self.bands[self.rank]), int(self.bands[self.rank + 1])
        self.fb[y0:y1].zero_()                                   # image_.clear() on the rows this rank renders
        self.dev.trace_frame(self.cam, self.shade, self.tiles, self.fb)   # this rank's band in one persistent launch
        self.comm.image_gather_rows(self.fb, self.W, self.bands, 0)

    def frame_host(self):
        self.frame_device()
        if self.rank == 0:
            self.fb_host.copy_(self.fb, non_blocking=True)
        self.dev.synchronize()

    def count_rays(self):
        """Deterministic per-frame ray counts of this rank's band (one extra frame with stats)."""
        tot = {k: 0 for k, _ in capi.Stats._fields_}   # a rank may own no tile at all
        self.fb.zero_()
        for t in self.tiles:
            st = self.dev.trace_tile(self.cam, self.shade, t, self.fb, stats=True)
            for k, v in st.as_dict().items():
                tot[k] = tot.get(k, 0) + v
        return tot


class InsituRenderer:
    def __init__(self, dev, scene, cam, shade, W, H, rank, world, comm, round_robin=False):
        import torch
        self.dev, self.cam, self.shade, self.W, self.H, self.rank, self.world, self.comm = dev, cam, shade, W, H, rank, world, comm
        wb = np.stack([d["bounds"] for d in scene["domains"]])
        self.domain_to_rank = capi.partition_domains(wb, scenes.scene_bounds(scene), world, round_robin)
        self.owned = [i for i, r in enumerate(self.domain_to_rank) if r == rank]
        dev.load_scene(scene, owned=self.owned)
        self.tiles = band_tiles(W, (0, H), shade.pixel_samples)   # every rank walks every blocking tile
        self.fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
        self.fb_host = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory() if rank == 0 else None

    def frame_device(self):
        self.fb.zero_()
        for t in self.tiles:
            self.comm.trace_tile_insitu(self.cam, self.shade, t, self.fb)
        self.comm.image_reduce(self.fb, self.W * self.H, 0)

    def frame_host(self):
        self.frame_device()
        if self.rank == 0:
            self.fb_host.copy_(self.fb, non_blocking=True)
        self.dev.synchronize()

    def count_rays(self):
        tot = {k: 0 for k, _ in capi.Stats._fields_}   # a rank may own no tile at all
        self.fb.zero_()
        for t in self.tiles:
            st = self.comm.trace_tile_insitu(self.cam, self.shade, t, self.fb, stats=True)
            for k, v in st.as_dict().items():
                tot[k] = tot.get(k, 0) + v
        return tot
