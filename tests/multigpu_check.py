"""Multi-rank GPU check, launched by torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tests/multigpu_check.py

Every rank renders with the in-situ domain partition and with the image partition through the C ABI's own
NCCL communicator; rank 0 compares both with a single-GPU render of the whole scene.  Used by
tests/test_gpu_multi.py when >= 2 GPUs are visible."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from spray_b200 import capi, multigpu, scenes
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.current_stream().cuda_stream
    sc = scenes.wavelets64()
    p = scenes.WAVELETS64_CAMERA
    W = H = 384
    cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, H)
    cam = capi.make_camera(cd)
    ok = True
    for kw in (dict(bounces=1, pixel_samples=1), dict(bounces=3, pixel_samples=2)):
        sh = capi.make_shade(lights=sc["lights"], **kw, **scenes.WAVELETS64_BLINN)
        # in-situ: domains sharded, NCCL min-all-reduce on the keys
        d1 = capi.Device(local, stream=stream)
        c1 = multigpu.bootstrap_comm(d1, rank, world)
        r1 = multigpu.InsituRenderer(d1, sc, cam, sh, W, H, rank, world, c1)
        assert len(r1.owned) == 64 // world
        st1 = r1.count_rays()
        r1.frame_device(); d1.synchronize()
        img_insitu = r1.fb.cpu().numpy()
        # image partition: all domains everywhere, row bands, NCCL gather
        d2 = capi.Device(local, stream=stream)
        c2 = multigpu.bootstrap_comm(d2, rank, world)
        r2 = multigpu.ImagePartitionRenderer(d2, sc, cam, sh, W, H, rank, world, c2, interleave=True)
        r2.frame_device(); d2.synchronize()
        img_image = r2.fb.cpu().numpy()
        d3 = capi.Device(local, stream=stream)
        c3 = multigpu.bootstrap_comm(d3, rank, world)
        r3 = multigpu.ImagePartitionRenderer(d3, sc, cam, sh, W, H, rank, world, c3, interleave=False)
        r3.frame_device(); d3.synchronize()
        img_rows = r3.fb.cpu().numpy()
        c3.close(); d3.close()
        # keys: local trace of the owned domains + min-all-reduce == the full single-device answer
        n = W * H
        rays = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
        d1.gen_eye_rays(cam, (0, 0, W, H), 1, rays)
        keys = torch.zeros(n, dtype=torch.int64, device="cuda")
        d1.intersect_domains(rays, n, keys, None)
        c1.allreduce_keys_min(keys, n)
        keys_full = torch.zeros(n, dtype=torch.int64, device="cuda")
        d2.intersect_domains(rays, n, keys_full, None)
        d1.synchronize(); d2.synchronize()
        if rank == 0:
            def render(tiles):
                fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
                tot = {}
                for t in tiles:
                    st = d2.trace_tile(cam, sh, t, fb, stats=True)
                    for k, v in st.as_dict().items():
                        tot[k] = tot.get(k, 0) + v
                d2.synchronize()
                return fb.cpu().numpy(), tot
            # sample ids (RNG seeds) are tile-relative, as in the reference (ooc_tracer.inl:84-172): compare each
            # decomposition with a single-GPU render over the SAME blocking tiles
            ref, tot = render(multigpu.band_tiles(W, (0, H), sh.pixel_samples))
            bb, bo = r2.block_begin, r2.block_owner          # interleaved row blocks (load balancing)
            ref_bands, _ = render([t for b in range(len(bo)) for t in multigpu.band_tiles(W, bb[b:b + 2], sh.pixel_samples)])
            # contiguous bands + row gather (the non-interleaved variant)
            bands = multigpu.row_bands(H, world)
            ref_rows, _ = render([t for r in range(world) for t in multigpu.band_tiles(W, bands[r:r + 2], sh.pixel_samples)])
            e3 = float(np.sqrt(np.mean((img_rows[..., :3].astype(np.float64) - ref_rows[..., :3]) ** 2)))
            e1 = float(np.sqrt(np.mean((img_insitu[..., :3].astype(np.float64) - ref[..., :3]) ** 2)))
            e2 = float(np.sqrt(np.mean((img_image[..., :3].astype(np.float64) - ref_bands[..., :3]) ** 2)))
            same_keys = bool(torch.equal(keys, keys_full))
            same_stats = all(st1[k] == tot[k] for k in ("eye_rays", "radiance_rays", "shadow_rays", "hits"))
            print("multigpu_check %s world=%d: insitu rmse %.3g | image-partition rmse %.3g | keys equal %s | stats equal %s (%s)"
                  % (kw, world, e1, e2, same_keys, same_stats, {k: st1[k] for k in ("radiance_rays", "shadow_rays", "hits")}))
            ok = ok and e1 <= 1e-6 and e2 <= 1e-6 and e3 <= 1e-6 and same_keys and same_stats and ref[..., :3].max() > 0.1
        c1.close(); c2.close(); d1.close(); d2.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, src=0)
    dist.destroy_process_group()
    if rank == 0:
        print("MULTIGPU_CHECK_OK" if ok else "MULTIGPU_CHECK_FAILED")
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
