"""One GPU renders the share that rank r of N would render under the image partition (interleaved row blocks), for every r:
how far does the per-rank frame shrink with N, and how unequal are the ranks?  (No communication: the flush goes nowhere.)
    python tools/share_one.py [workload]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    import bench
    from spray_b200 import capi, multigpu, scenes
    wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream()
    sc, cd, shade, label, W, H = bench.workload(wl)
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **shade)
    dev = capi.Device(0, stream=stream.cuda_stream)
    dev.load_scene(sc)
    fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    blocking = scenes.image_schedule_tiles(W, H, sh.pixel_samples, 1, 0)
    warm = os.environ.get("SHARE_WARM", "0") == "1"       # 1: no L2 flush between frames
    for world in (1, 2, 4, 8, 16, 32, 64):
        bb, bo = multigpu.interleaved_blocks(H, world)
        times, rays = [], []
        for rank in range(min(world, 8)):
            rects, parents = multigpu.rects_of_blocks(bb, bo, rank, blocking)
            st = dev.trace_frame_rects(cam, sh, rects, parents, fb, stats=True)
            rays.append(st.radiance_rays + st.shadow_rays)
            ms = []
            for it in range(8):
                if not warm:
                    flush.fill_(1)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                dev.trace_frame_rects(cam, sh, rects, parents, fb)
                b.record(stream)
                dev.synchronize()
                if it >= 3:
                    ms.append(a.elapsed_time(b))
            times.append(float(np.mean(ms)))
        print("share_one %s%s N=%d: %d blocks of %d rows | per-rank ms %s | max %.3f mean %.3f | rays per rank min %d max %d"
              % (wl, " warm" if warm else "", world, len(bo), bb[1] - bb[0], " ".join("%.3f" % t for t in times), max(times), float(np.mean(times)), min(rays), max(rays)), flush=True)
    dev.close()


if __name__ == "__main__":
    main()
