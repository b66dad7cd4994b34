#!/usr/bin/env python3
"""In-situ wavefront tile on ONE rank (communicator of size 1): kernel-time breakdown under
`ncu --metrics gpu__time_duration.sum` without the multi-rank launch."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from spray_b200 import capi, scenes, multigpu

sc = scenes.wavelets64()
p = scenes.WAVELETS64_CAMERA
W = H = 2048
cd = scenes.camera_frame(p["pos"], p["lookat"], p["up"], p["vfov"], W, H)
cam = capi.make_camera(cd)
sh = capi.make_shade(lights=sc["lights"], bounces=1, pixel_samples=1, **scenes.WAVELETS64_BLINN)
dev = capi.Device(0, stream=torch.cuda.current_stream().cuda_stream)
comm = capi.Comm(dev, 1, 0, capi.comm_unique_id())
owned = list(range(0, 64, int(os.environ.get("OWN_STRIDE", "1"))))
dev.load_scene(sc, owned=owned)
fb = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
tiles = multigpu.band_tiles(W, (0, H), 1)
for it in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    fb.zero_()
    for t in tiles:
        comm.trace_tile_insitu(cam, sh, t, fb)
    torch.cuda.synchronize(); print("frame ms %.3f (owned %d domains)" % ((time.perf_counter() - t0) * 1e3, len(owned)))
comm.close(); dev.close()
