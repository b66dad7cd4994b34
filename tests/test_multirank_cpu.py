"""World-size-2 checks of the multi-GPU HOST logic on CPU (gloo): the domain -> rank partition, the
key / occlusion resolution rule of the in-situ path (min-all-reduce on packed 64-bit keys replaces
insitu::VBuf::compositeTbuf's MPI_Allreduce(MIN), max-all-reduce replaces compositeObuf's BOR;
reference src/insitu/insitu_vbuf.h:109-129) and the image-partition row bands + gather
(reference src/render/tile.cc:317-391, src/display/image.h:167-181).

The per-rank "device" here is the CPU oracle restricted to the rank's domains — test infrastructure standing in
for libspraygpu so that the sharding rules can run without a GPU; the same protocol runs on NCCL in
tests/test_gpu_multi.py / tests/multigpu_check.py."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

I64_MAX = np.iinfo(np.int64).max


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _keys_to_i64(k):
    """uint64 keys -> int64 preserving order: every real key has bit 63 clear (t > 0), MISS maps to INT64_MAX."""
    k = np.asarray(k, np.uint64)
    out = k.astype(np.int64)
    out[k == np.uint64(0xFFFFFFFFFFFFFFFF)] = I64_MAX
    return out


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from oracle import pyoracle as O
    from spray_b200 import multigpu, scenes
    import partition_ref as partition
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        sc = scenes.wavelets64()
        nd = len(sc["domains"])
        osc = O.Scene(sc["domains"])
        p = scenes.WAVELETS64_CAMERA
        W = H = 64
        cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, H)
        cam = O.camera_from_desc(cd)
        rays = O.gen_eye_rays(cam, (0, 0, W, H))
        n = len(rays)

        # ---- in-situ: Morton partition, every rank traces every ray against ITS domains
        wb = np.stack([d["bounds"] for d in sc["domains"]])
        d2r = partition.insitu_partition(wb, scenes.scene_bounds(sc), world)
        owned = [i for i in range(nd) if d2r[i] == rank]
        assert len(owned) == nd // world
        counts, ids, tmins = osc.domain_lists(rays, cap=nd)
        entered = np.zeros((n, nd), bool)
        for i in range(n):
            entered[i, ids[i, :counts[i]]] = True
        local = np.full(n, I64_MAX, np.int64)
        occ_local = np.zeros(n, np.uint8)
        for d in owned:
            kd, _ = osc.intersect_domain(d, rays)
            kd = _keys_to_i64(kd)
            kd[~entered[:, d]] = I64_MAX
            local = np.minimum(local, kd)
            od = np.asarray(osc.occluded_domain(d, rays), np.uint8)
            od[~entered[:, d]] = 0
            occ_local = np.maximum(occ_local, od)
        tk = torch.from_numpy(local.copy())
        dist.all_reduce(tk, op=dist.ReduceOp.MIN)
        to = torch.from_numpy(occ_local.copy())
        dist.all_reduce(to, op=dist.ReduceOp.MAX)
        k_full, _ = osc.intersect(rays)
        ok_keys = bool(np.array_equal(tk.numpy(), _keys_to_i64(k_full)))
        ok_occ = bool(np.array_equal(to.numpy(), np.asarray(osc.occluded(rays), np.uint8)))
        # the winner's domain field names exactly one owner rank
        hit = tk.numpy() != I64_MAX
        dom = ((tk.numpy()[hit] >> 23) & 0x1FF).astype(np.int64)
        mine = int(np.sum(d2r[dom] == rank))
        tm = torch.tensor([mine], dtype=torch.int64)
        dist.all_reduce(tm)
        ok_owner = int(tm.item()) == int(hit.sum())

        # ---- in-situ with ray forwarding (insitu3.cu's protocol): a sample lives on its HOME rank (interleaved row blocks);
        # its ray is sent to every rank that owns an entered domain, flagged primary for the owner of the NEAREST one;
        # pass 0 traces the primary records, pass 1 the others and keeps only what beats the known key; the home rank's
        # key array ends up with the single-rank closest hit and every hit has exactly one attribute writer
        rpb = 4
        home = ((np.arange(n) // W) // rpb) % world                  # pixel row -> home rank (1 spp, whole-image tile)
        first = ids[np.arange(n), 0]                                  # nearest entered domain (DomainList::sort order)
        outbox = [[] for _ in range(world)]
        for i in np.nonzero(home == rank)[0]:
            if counts[i] == 0:
                continue
            owners = set(int(d2r[d]) for d in ids[i, :counts[i]])
            for r in owners:
                outbox[r].append((int(i), int(d2r[first[i]]) == r))
        gathered_boxes = [None] * world
        dist.all_gather_object(gathered_boxes, outbox)
        inbox = [rec for src in gathered_boxes for rec in src[rank]]           # what every home rank sent to ME
        dom_keys = {d: _keys_to_i64(osc.intersect_domain(d, rays)[0]) for d in owned}
        for d in owned:
            dom_keys[d][~entered[:, d]] = I64_MAX
        home_keys = torch.full((n,), I64_MAX, dtype=torch.int64)
        writers = torch.zeros(n, dtype=torch.int64)
        for want_primary in (True, False):
            mine_k = np.full(n, I64_MAX, np.int64)
            known = home_keys.numpy().copy()                                     # after the barrier: pass 0's result
            for (i, primary) in inbox:
                if primary != want_primary:
                    continue
                k = min(int(dom_keys[d][i]) for d in owned)
                if k < known[i]:                                                 # tbound: only hits below the known key
                    mine_k[i] = k
            t = torch.from_numpy(np.minimum(mine_k, home_keys.numpy()))
            dist.all_reduce(t, op=dist.ReduceOp.MIN)                             # = atomicMin_system into the home arrays
            home_keys = t
        final = home_keys.numpy()
        for (i, primary) in inbox:                                               # F3: candidates whose key survived
            if final[i] != I64_MAX and min(int(dom_keys[d][i]) for d in owned) == final[i]:
                writers[i] += 1
        dist.all_reduce(writers)
        ok_fwd = bool(np.array_equal(final, _keys_to_i64(k_full))) and bool(np.array_equal(writers.numpy() == 1, final != I64_MAX))
        ok_keys = ok_keys and ok_fwd

        # ---- image partition: row bands, each rank renders its band, rank 0 gathers
        bands = multigpu.row_bands(H, world)
        tiles = multigpu.band_tiles(W, bands[rank:rank + 2], 1, max_samples=W * 8)
        assert sum(t[3] for t in tiles) == bands[rank + 1] - bands[rank]
        sh, keep = O.make_shade(bounces=2, pixel_samples=1, lights=sc["lights"], **scenes.WAVELETS64_BLINN)
        fb, st = osc.render(cam, sh, tiles)
        band = torch.from_numpy(np.ascontiguousarray(fb[bands[rank]:bands[rank + 1]]).astype(np.float32))
        gathered = [torch.zeros((int(bands[r + 1] - bands[r]), W, 4), dtype=torch.float32) for r in range(world)] if rank == 0 else None
        dist.gather(band, gathered, dst=0)
        ok_img = True
        if rank == 0:
            img = torch.cat(gathered).numpy()
            # samid (the RNG seed of bounce rays) is the index within the blocking tile (ooc_tracer.inl:84-172), so
            # the single-rank reference frame must walk the same tiles
            all_tiles = [t for r in range(world) for t in multigpu.band_tiles(W, bands[r:r + 2], 1, max_samples=W * 8)]
            fb_full, st_full = osc.render(cam, sh, all_tiles)
            ok_img = bool(np.array_equal(img, fb_full.astype(np.float32)))
        q.put((rank, ok_keys, ok_occ, ok_owner, ok_img, len(owned)))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_partition_keys_occlusion_and_bands():
    import torch.multiprocessing as mp
    from oracle import pyoracle
    pyoracle.build()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_keys, ok_occ, ok_owner, ok_img, nown in sorted(res):
        assert ok_keys, "rank %d: min-all-reduce of per-rank keys != single-rank closest hit" % rank
        assert ok_occ, "rank %d: max-all-reduce of per-rank occlusion != single-rank any-hit" % rank
        assert ok_owner, "rank %d: winner ownership is not a partition of the hits" % rank
        assert ok_img, "rank %d: gathered row bands != full frame" % rank
        assert nown == 32


def test_row_bands_and_band_tiles_cover_the_image():
    from spray_b200 import multigpu
    for H, world in ((2048, 8), (2048, 3), (7, 8), (1, 2)):
        b = multigpu.row_bands(H, world)
        assert b[0] == 0 and b[-1] == H and np.all(np.diff(b) >= 0) and np.diff(b).max() - np.diff(b).min() <= 1
        rows = 0
        for r in range(world):
            for (x, y, w, h) in multigpu.band_tiles(2048, b[r:r + 2], 4):
                assert x == 0 and w == 2048 and w * h * 4 <= 1024 * 1024 and b[r] <= y and y + h <= b[r + 1]
                rows += h
        assert rows == H


def test_interleaved_blocks_cover_and_balance():
    from spray_b200 import multigpu
    for H, world in ((2048, 8), (2048, 4), (4096, 8), (150, 3), (8, 8)):
        for share in (1.0, 0.5, 0.0):
            b, o = multigpu.interleaved_blocks(H, world, root=0, root_share=share)
            assert b[0] == 0 and b[-1] == H and np.all(np.diff(b) > 0) and len(o) == len(b) - 1
            assert np.all(np.diff(b)[:-1] % 4 == 0)                    # whole 8x4 pixel blocks except the ragged last one
            assert o.min() >= 0 and o.max() < world
            cnt = np.bincount(o, minlength=world)
            if share == 1.0:
                assert cnt.max() - cnt.min() <= 1
            if share == 0.0 and world > 1:
                assert cnt[0] == 0                                   # a pure compositor renders nothing
            if share < 1.0:
                assert cnt[0] <= cnt[1:].max()


def test_rects_of_blocks_partition_the_frame_and_keep_blocking_tiles():
    """Image partition v2: every rank renders rectangles of the ONE-rank blocking tiles (spraygpu_trace_frame_rects), so
    sample ids do not depend on the rank count.  The rectangles of all ranks must tile the image exactly, lie inside
    their blocking tile, start on 4-row boundaries (8x4 pixel blocks) and fit one persistent launch per rank."""
    from spray_b200 import multigpu, scenes
    for (W, H, spp, world) in ((2048, 2048, 1, 8), (2048, 2048, 1, 2), (4096, 4096, 4, 8), (4096, 4096, 4, 4), (320, 200, 3, 3)):
        blocking = scenes.image_schedule_tiles(W, H, spp, 1, 0)
        bb, bo = multigpu.interleaved_blocks(H, world)
        cover = np.zeros(H, np.int32)
        for rank in range(world):
            rects, parents = multigpu.rects_of_blocks(bb, bo, rank, blocking)
            assert len(rects) == len(parents)
            if (W, H) != (320, 200):
                assert len(rects) <= 32
            for (x, y, w, h), (px, py, pw, ph) in zip(rects, parents):
                assert x == px == 0 and w == pw == W and py <= y and y + h <= py + ph and h > 0
                assert (px, py, pw, ph) in blocking
                cover[y:y + h] += 1
        assert np.all(cover == 1)


def test_scatter_partition_is_a_balanced_deterministic_deal():
    """bench.py's in-situ deal (multigpu.scatter_partition; not a reference partition): every rank gets nd / N domains,
    the deal is a pure function of the domain boxes, and on the 4 x 4 x 4 grid it spreads the camera-facing cells where
    contiguous Morton runs and the round-robin deal leave one rank without any."""
    from spray_b200 import capi, multigpu, scenes
    sc = scenes.wavelets64()
    wb = np.stack([d["bounds"] for d in sc["domains"]])
    sb = scenes.scene_bounds(sc)
    for world in (2, 4, 8):
        a = multigpu.scatter_partition(wb, sb, world)
        assert np.array_equal(a, multigpu.scatter_partition(wb, sb, world))
        assert np.array_equal(np.bincount(a, minlength=world), np.full(world, len(wb) // world))
    lo = wb[:, :3]
    step = np.diff(np.unique(lo[:, 0]))[0]
    idx = np.round((lo - lo.min(0)) / step).astype(int)
    front = (idx == idx.max()).any(1)
    rr = capi.partition_domains(wb, sb, 8, True)
    sca = multigpu.scatter_partition(wb, sb, 8)
    assert np.bincount(rr[front], minlength=8).min() == 0            # the parity partition starves the all-even rank
    assert np.bincount(sca[front], minlength=8).min() >= 2
    assert np.bincount(sca[front], minlength=8).max() <= np.bincount(rr[front], minlength=8).max()
