#!/usr/bin/env python3
"""Per-source-line summary of an ncu report's source page (needs -lineinfo + --import-source on).

    ncu -i prof.ncu-rep --page source --csv --print-source cuda,sass > src.csv
    tools/ncu_lines.py src.csv [top N]

Prints, for the hottest source lines: samples, warp instructions executed, average active threads per
instruction (SIMD efficiency) and the dominant stall reasons."""
import csv
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    rows = list(csv.reader(open(path)))
    cur_file = None
    hdr = None
    out = []
    tot_inst = tot_tinst = tot_samples = 0
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or r[0] == "":
            continue
        d = dict(zip(hdr, r))
        # the two "Source" columns collide in the dict; take by position
        line_no, src = r[0], r[1]
        try:
            samples = int(r[hdr.index("# Samples")])
            inst = int(r[hdr.index("Instructions Executed")])
            tinst = int(r[hdr.index("Thread Instructions Executed")])
        except ValueError:
            continue
        stalls = {}
        for k in hdr:
            if k.startswith("stall_") and "Not Issued" not in k:
                try:
                    stalls[k[6:]] = int(d[k])
                except ValueError:
                    pass
        out.append((samples, inst, tinst, cur_file, line_no, src.strip()[:90], stalls))
        tot_inst += inst
        tot_tinst += tinst
        tot_samples += samples
    print("total: samples %d  warp-inst %d  threads/inst %.2f" % (tot_samples, tot_inst, tot_tinst / max(tot_inst, 1)))
    out.sort(key=lambda x: -x[1])
    print("%6s %6s %5s  %-14s %5s  %s" % ("inst%", "smpl%", "thr", "file", "line", "source | top stalls"))
    for samples, inst, tinst, f, ln, src, stalls in out[:top]:
        st = sorted(stalls.items(), key=lambda kv: -kv[1])[:3]
        sts = " ".join("%s:%d" % (k, v) for k, v in st if v)
        print("%6.2f %6.2f %5.1f  %-14s %5s  %s | %s" % (100.0 * inst / max(tot_inst, 1), 100.0 * samples / max(tot_samples, 1),
                                                        tinst / max(inst, 1), f, ln, src, sts))


if __name__ == "__main__":
    main()
