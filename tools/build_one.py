#!/usr/bin/env python3
"""Commit ONE soup domain of N triangles a few times (kernel breakdown of spraygpu_scene_commit under ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spray_b200 import capi, scenes
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
sc = scenes.soup_scene((1, 1, 1), n)
dev = capi.Device(0)
d = sc["domains"][0]
dev.set_domains([d["bounds"]], [d.get("mtl", (0, (1, 1, 1)))])
s = dev.new_scene()
s.set_mesh(d["verts"], d["faces"], d.get("rgb"))
for it in range(3):
    s.set_transform(None)
    s.commit()
    print("commit %d tris: %.3f ms device, %d nodes" % (n, s.info().build_ms, s.info().num_nodes))
dev.close()
