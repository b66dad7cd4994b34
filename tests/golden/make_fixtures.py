#!/usr/bin/env python3
"""Generate the committed golden fixtures from the reference's example inputs.

Run ONCE in the build container (where /root/reference is mounted read-only):

    python tests/golden/make_fixtures.py

Outputs (committed, small):
  tests/golden/wavelet_mesh.npz      geometry of examples/wavelet.ply (config C1/C2 input), decoded
                                      with an independent numpy PLY reader: verts f32[2840,3],
                                      rgb u8[2840,3], faces u32[5480,3]
  tests/golden/wavelets64_scene.json the domain stanzas of examples/wavelets64/wavelets64.spray
                                      (bounds / translate / counts / light) for checking our own
                                      .spray writer + loader against the reference's file

Nothing under tests/, bench.py or smoke() reads /root/reference at run time; they read these files.
The reference repository holds no tests, golden vectors or images usable as numeric goldens
(SURVEY.md section 4), so these INPUT fixtures are all the reference can pin.
"""
import json
import os
import sys

import numpy as np

REF = "/root/reference/examples"
HERE = os.path.dirname(os.path.abspath(__file__))


def read_ply_binary_le(path):
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"end_header\n") + len(b"end_header\n")
    header = data[:end].decode("ascii").splitlines()
    assert header[0] == "ply" and header[1].startswith("format binary_little_endian")
    nv = nf = None
    vprops = []
    section = None
    for line in header:
        tok = line.split()
        if tok[:2] == ["element", "vertex"]:
            nv = int(tok[2]); section = "v"
        elif tok[:2] == ["element", "face"]:
            nf = int(tok[2]); section = "f"
        elif tok[0] == "property" and section == "v":
            vprops.append((tok[2], tok[1]))
    assert [p[0] for p in vprops] == ["x", "y", "z", "red", "green", "blue"], vprops
    vdt = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("r", "u1"), ("g", "u1"), ("b", "u1")])
    v = np.frombuffer(data, dtype=vdt, count=nv, offset=end)
    off = end + nv * vdt.itemsize
    fdt = np.dtype([("n", "u1"), ("i", "<i4", (3,))])
    fc = np.frombuffer(data, dtype=fdt, count=nf, offset=off)
    assert off + nf * fdt.itemsize == len(data)
    assert (fc["n"] == 3).all()
    verts = np.stack([v["x"], v["y"], v["z"]], axis=1).astype(np.float32)
    rgb = np.stack([v["r"], v["g"], v["b"]], axis=1).astype(np.uint8)
    faces = fc["i"].astype(np.uint32)
    return verts, rgb, faces


def parse_spray(path):
    doms, lights = [], []
    cur = None
    for line in open(path):
        tok = line.split()
        if not tok or tok[0].startswith("#"):
            continue
        if tok[0] == "domain":
            cur = {"translate": [0.0, 0.0, 0.0]}
            doms.append(cur)
        elif tok[0] == "light":
            lights.append({"type": tok[1], "v": [float(x) for x in tok[2:]]})
        elif tok[0] == "file":
            cur["file"] = tok[1]
        elif tok[0] == "mtl":
            cur["mtl"] = [tok[1]] + [float(x) for x in tok[2:]]
        elif tok[0] == "bound":
            cur["bound"] = [float(x) for x in tok[1:]]
        elif tok[0] in ("face", "vertex"):
            cur[tok[0]] = int(tok[1])
        elif tok[0] == "translate":
            cur["translate"] = [float(x) for x in tok[1:]]
        else:
            raise SystemExit("unhandled tag " + tok[0])
    return {"domains": doms, "lights": lights}


def main():
    if not os.path.isdir(REF):
        sys.exit("reference not mounted; fixtures are already committed")
    verts, rgb, faces = read_ply_binary_le(os.path.join(REF, "wavelet.ply"))
    np.savez_compressed(os.path.join(HERE, "wavelet_mesh.npz"), verts=verts, rgb=rgb, faces=faces)
    scene = parse_spray(os.path.join(REF, "wavelets64", "wavelets64.spray"))
    with open(os.path.join(HERE, "wavelets64_scene.json"), "w") as f:
        json.dump(scene, f, indent=0)
    print("verts", verts.shape, "faces", faces.shape, "bbox", verts.min(0), verts.max(0),
          "domains", len(scene["domains"]))


if __name__ == "__main__":
    main()
