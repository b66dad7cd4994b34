"""One rank, whole scene: the wavefront in-situ tile (peer-memory and NCCL variants) against the fused kernel on a
bench workload — the per-kernel breakdown of the in-situ path (run under `ncu --metrics gpu__time_duration.sum`).
    python tools/insitu_one.py [workload] [frames]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import bench
    from spray_b200 import capi, multigpu
    wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream()
    sc, cd, shade, label, W, H = bench.workload(wl)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **shade)
    dev = capi.Device(0, stream=stream.cuda_stream)
    comm = capi.Comm(dev, 1, 0, capi.comm_unique_id())
    for mode in (sys.argv[3].split(",") if len(sys.argv) > 3 else ("p2p", "nccl")):
        r = multigpu.InsituRenderer(dev, sc, cam, sh, W, H, 0, 1, comm, mode=mode)
        st = r.count_rays()
        rays = st["radiance_rays"] + st["shadow_rays"]
        for _ in range(2):
            r.frame_device()
        dev.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = dev.launch_count()
        e0.record(stream)
        for _ in range(frames):
            r.frame_device()
        e1.record(stream)
        dev.synchronize()
        ms = e0.elapsed_time(e1) / frames
        print("insitu_one %s %s: %d tiles, %.3f ms/frame, %.0f Mrays/s, %d launches/frame" %
              (wl, mode, len(r.tiles), ms, rays / ms / 1e3, (dev.launch_count() - l0) // frames))
        r.close()
    comm.close(); dev.close()


if __name__ == "__main__":
    main()
