"""How much of the host-image flush hides behind the next frame's trace?  K frames of C2, wall bracket (events on the
device stream + final host wait), no L2 fill:  trace only | trace + synchronous flush | trace + flush_async."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import bench
    from spray_b200 import capi, scenes
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream()
    sc, cd, shade, label, W, H = bench.workload("c2")
    cam = capi.make_camera(cd)
    sh = capi.make_shade(lights=sc["lights"], **shade)
    dev = capi.Device(0, stream=stream.cuda_stream)
    dev.load_scene(sc)
    tiles = scenes.image_schedule_tiles(W, H, 1, 1, 0)
    img = capi.Image(dev, None, W, H)
    K = 40

    def run(name, step, finish):
        for _ in range(5):
            step()
        finish(); dev.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record(stream)
        for _ in range(K):
            step()
        finish()
        b.record(stream)
        dev.synchronize()
        print("pipe_probe %-28s %.3f ms/frame (events) %.3f ms/frame (wall)" % (name, a.elapsed_time(b) / K, (time.perf_counter() - t0) * 1e3 / K), flush=True)

    def s_trace():
        dev.trace_frame(cam, sh, tiles, img.device)
        img.flush(clear=True)

    def s_sync():
        dev.trace_frame(cam, sh, tiles, img.device)
        img.flush(to_host=True, clear=True)

    def s_async():
        dev.trace_frame(cam, sh, tiles, img.device)
        img.flush_async(to_host=True)

    evs = []

    def s_async_ev():
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        dev.trace_frame(cam, sh, tiles, img.device)
        b.record(stream)
        evs.append((a, b))
        img.flush_async(to_host=True)

    def s_async_clear():
        dev.trace_frame(cam, sh, tiles, img.device)
        img.flush_async(to_host=False)

    run("trace + clear only", s_trace, lambda: None)
    run("trace + sync flush to host", s_sync, lambda: None)
    run("trace + flush_async", s_async, img.wait)
    run("trace + flush_async (no host)", s_async_clear, img.wait)
    run("trace + flush_async, events", s_async_ev, img.wait)
    dur = [a.elapsed_time(b) for a, b in evs[-K:]]
    gaps = [evs[i][1].elapsed_time(evs[i + 1][0]) for i in range(len(evs) - K, len(evs) - 1)]
    print("pipe_probe   trace kernel inside the pipelined loop: mean %.3f ms (min %.3f max %.3f); gap between traces mean %.3f ms"
          % (sum(dur) / len(dur), min(dur), max(dur), sum(gaps) / len(gaps)), flush=True)
    img.close(); dev.close()


if __name__ == "__main__":
    main()
