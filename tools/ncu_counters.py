#!/usr/bin/env python3
"""Turn an ncu --set full capture of render_tile_kernel into the entry bench.py prints beside its timings.

    ncu -i gpurun_out/prof_c2.ncu-rep --page raw --csv > gpurun_out/prof_c2_raw.csv      (here, no GPU needed)
    tools/ncu_counters.py <workload> <raw.csv> <lib.so that was profiled> [source note]

Writes / updates profiles/ncu_counters.json:  { build id: { workload: {counters...} } }, build id = capi.build_id() =
sha256 of the sources the library is built from (run this on the SAME tree that was profiled).  bench.py looks its
own build id up in that file and prints `traffic: null` when there is no capture of that exact build, so a
profile can never be quoted beside another build's timings (the traffic constant was two kernel
revisions stale).  Values are per LAUNCH (one launch = one frame); the kernel row with the longest duration is used."""
import csv
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "ncu_counters.json")

# output key -> candidate raw-page metric names (first present wins)
WANT = {
    "dram_bytes_read": ["dram__bytes_read.sum"],
    "dram_bytes_write": ["dram__bytes_write.sum"],
    "l2_bytes": ["lts__t_bytes.sum"],
    "l2_sectors": ["lts__t_sectors.sum"],
    "l1_throughput_pct": ["l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"],
    "l2_throughput_pct": ["lts__throughput.avg.pct_of_peak_sustained_elapsed"],
    "dram_throughput_pct": ["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed"],
    "issue_slots_busy_pct": ["sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct"],
    "lanes_per_inst": ["smsp__thread_inst_executed_per_inst_executed.ratio"],
    "l1_hit_pct": ["l1tex__t_sector_hit_rate.pct"],
    "l2_hit_pct": ["lts__t_sector_hit_rate.pct"],
    "registers": ["launch__registers_per_thread"],
    "achieved_warps_per_sm": ["sm__warps_active.avg.per_cycle_active"],
    "eligible_warps_per_scheduler": ["smsp__warps_eligible.avg.per_cycle_active"],
    "ncu_duration_ns": ["gpu__time_duration.sum"],
    "inst_executed": ["smsp__inst_executed.sum", "sm__inst_executed.sum"],
}


def num(x):
    try:
        return float(str(x).replace(",", ""))
    except Exception:
        return None


def main():
    wl, raw, lib = sys.argv[1], sys.argv[2], sys.argv[3]
    note = sys.argv[4] if len(sys.argv) > 4 else os.path.basename(raw)
    rows = list(csv.reader(open(raw)))
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names, units = rows[hdr], rows[hdr + 1]
    data = [r for r in rows[hdr + 2:] if len(r) == len(names)]
    kcol = names.index("Kernel Name")
    data = [r for r in data if "render_tile" in r[kcol]]
    if not data:
        raise SystemExit("no render_tile kernel row in " + raw)
    dcol = names.index("gpu__time_duration.sum") if "gpu__time_duration.sum" in names else None
    row = max(data, key=lambda r: num(r[dcol]) or 0.0) if dcol is not None else data[-1]
    ent = {"kernel": row[kcol], "source": note}
    for key, cands in WANT.items():
        for c in cands:
            if c in names:
                v = num(row[names.index(c)])
                u = units[names.index(c)]
                if v is not None:
                    # ncu scales units on the raw page (Kbyte, Mbyte, usecond ...): normalise to bytes / ns
                    scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "byte": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6,
                             "msecond": 1e6, "ns": 1.0, "nsecond": 1.0, "s": 1e9, "second": 1e9}.get(u, 1.0)
                    ent[key] = v * scale if key in ("dram_bytes_read", "dram_bytes_write", "l2_bytes", "ncu_duration_ns") else v
                    break
    if "l2_bytes" not in ent and "l2_sectors" in ent:
        ent["l2_bytes"] = ent["l2_sectors"] * 32.0
    if "ncu_duration_ns" in ent:
        ent["ncu_duration_ms"] = ent.pop("ncu_duration_ns") / 1e6
    sys.path.insert(0, ROOT)
    from spray_b200 import capi
    h = capi.build_id()        # sha256 of the library's sources (the binary is not bit-reproducible); `lib` is kept for the log
    db = {}
    if os.path.exists(OUT):
        db = json.load(open(OUT))
    db.setdefault(h, {})[wl] = ent
    json.dump(db, open(OUT, "w"), indent=1, sort_keys=True)
    print(h[:16], wl, json.dumps(ent))


if __name__ == "__main__":
    main()
