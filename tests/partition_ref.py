"""Domain -> GPU map: the reference's in-situ Morton partition (src/render/data_partition.h:59-137,
src/render/morton.h:33-52), restated in numpy float32.  TEST INFRASTRUCTURE: an independent restatement that tests compare with the product's
spraygpu_partition_domains (comm.cu) and the oracle's orc_partition; the product never imports it."""
import numpy as np

f32 = np.float32


def _expand_bits(v):
    v = np.uint32(v)
    v = (v * np.uint32(0x00010001)) & np.uint32(0xFF0000FF)
    v = (v * np.uint32(0x00000101)) & np.uint32(0x0F00F00F)
    v = (v * np.uint32(0x00000011)) & np.uint32(0xC30C30C3)
    v = (v * np.uint32(0x00000005)) & np.uint32(0x49249249)
    return v


def morton30(x, y, z):
    q = [np.uint32(min(max(f32(c) * f32(1024.0), f32(0.0)), f32(1023.0))) for c in (x, y, z)]
    with np.errstate(over="ignore"):
        return int(_expand_bits(q[0]) * np.uint32(4) + _expand_bits(q[1]) * np.uint32(2) + _expand_bits(q[2]))


def morton_codes(world_bounds, scene_bounds):
    wb = np.asarray(world_bounds, np.float32).reshape(-1, 6)
    sb = np.asarray(scene_bounds, np.float32)
    diag = (sb[3:] - sb[:3]).astype(np.float32)
    sf = (f32(1.0) / diag).astype(np.float32)
    off = (f32(0.0) - (sb[:3] * sf).astype(np.float32)).astype(np.float32)
    codes = []
    for b in wb:
        c = np.array([f32(0.5 * float(f32(b[k] + b[3 + k]))) for k in range(3)], np.float32)
        p = ((sf * c).astype(np.float32) + off).astype(np.float32)
        codes.append(morton30(p[0], p[1], p[2]))
    return np.array(codes, np.uint32)


def insitu_partition(world_bounds, scene_bounds, num_ranks, round_robin=False):
    """domain id -> rank.  Contiguous Morton runs of ndomains/num_ranks (GROUP_CLOSE_DOMAINS branch,
    data_partition.h:117-137); round_robin=True is the reference's `#else` branch (:139-155) which
    scatters neighbouring domains for view-independent load balance."""
    codes = morton_codes(world_bounds, scene_bounds)
    nd = len(codes)
    order = sorted(range(nd), key=lambda i: (int(codes[i]), i))  # reference: unstable sort; ties by id here
    d2r = np.zeros(nd, np.int32)
    if round_robin:
        for k, dom in enumerate(order):
            d2r[dom] = k % num_ranks
        return d2r
    shares = max(nd // num_ranks, 1)
    rank = s = 0
    for dom in order:
        d2r[dom] = rank
        s += 1
        if s == shares:
            s = 0
            rank += 1
            if rank == num_ranks:
                rank = 0
    return d2r
