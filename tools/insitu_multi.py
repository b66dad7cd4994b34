"""Per-phase breakdown of the in-situ tiles at N ranks (torchrun): SPRAY_INSITU_TIMING=1 on one frame per mode.
    python -m torch.distributed.run --nproc-per-node N ... tools/insitu_multi.py [workload] [modes, comma-separated]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import bench
    from spray_b200 import capi, multigpu
    wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
    modes = (sys.argv[2] if len(sys.argv) > 2 else "p2p,fwd").split(",")
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.current_stream()
    sc, cd, shade, label, W, H = bench.workload(wl)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **shade)
    dev = capi.Device(local, stream=stream.cuda_stream)
    comm = multigpu.bootstrap_comm(dev, rank, world)
    for mode in modes:
        r = multigpu.InsituRenderer(dev, sc, cam, sh, W, H, rank, world, comm, mode=mode, round_robin=os.environ.get("SPRAY_INSITU_RR", "0") == "1", scatter=os.environ.get("SPRAY_INSITU_PART", "") == "scatter", shm_name="spraygpu_tm_%d_%s" % (os.getppid(), mode))
        st = r.count_rays()
        rays = st["radiance_rays"] + st["shadow_rays"]
        for _ in range(3):
            r.frame_device()
        dev.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(5):
            r.frame_device()
        e1.record(stream)
        dev.synchronize()
        ms = e0.elapsed_time(e1) / 5
        if rank == 0:
            print("insitu_multi rr=%s %s %s x%d: %d tiles, %.3f ms/frame, %.0f Mrays/s" % (os.environ.get("SPRAY_INSITU_RR", "0"), wl, mode, world, len(r.tiles), ms, rays / ms / 1e3), flush=True)
        dist.barrier()
        os.environ["SPRAY_INSITU_TIMING"] = "1"
        r.frame_device()
        dev.synchronize()
        os.environ["SPRAY_INSITU_TIMING"] = "0"
        dist.barrier()
        r.close()
    comm.close(); dev.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
