#!/usr/bin/env python
"""Turn the ncu reports a gpurun call brought back (gpurun_out/*.ncu-rep, *_launches.csv) into the small, tracked
summaries under profiles/r2/.  usage: python tools/make_profiles.py"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "r2")
GO = os.path.join(ROOT, "gpurun_out")
KEYS = {"gpu__time_duration.sum": "time", "smsp__inst_executed.sum": "warp_inst", "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct", "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct", "launch__registers_per_thread": "regs", "launch__grid_size": "grid",
        "lts__t_sector_hit_rate.pct": "l2_hit_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio": "stall_long_sb", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio": "stall_wait",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio": "stall_barrier", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio": "stall_short_sb"}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0}


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {"kernel": vals[hdr.index("Kernel Name")].split("(")[0], "grid": vals[hdr.index("Grid Size")], "block": vals[hdr.index("Block Size")]}
    for k, name in KEYS.items():
        if k in hdr:
            v, u = float(vals[hdr.index(k)].replace(",", "")), units[hdr.index(k)]
            d[name] = v * UNIT[u] if u in UNIT else v
    return d


def main():
    os.makedirs(OUT, exist_ok=True)
    caps = {"k_fast_out int64, config 3 (50k nodes x 100k pods)": "r2f_out64", "k_fast_out uint32 scores (YODA_EVAL_RAW32), config 3": "r2f_out32",
            "k_fast_out int64, config 4 per-GPU shape at N=8 (125k nodes x 10k pods)": "r2f_out64_c4", "k_fast_score, config 3": "r2f_score",
            "k_fast_score, 125k nodes x 10k pods": "r2f_score_c4"}
    doc = {"how": "ncu --set full --import-source on --clock-control none -k regex:<kernel> -c 1 python bench.py --steps 1 --warmup 1 [shape]; one capture per kernel, "
                  "times are cold-cache and serialised", "captures": {}}
    for what, name in caps.items():
        rep = os.path.join(GO, name + ".ncu-rep")
        if os.path.exists(rep):
            doc["captures"][what] = raw(rep)
    with open(os.path.join(OUT, "ncu_full_hot_kernels.json"), "w") as f:
        json.dump(doc, f, indent=1)
    o = doc["captures"].get("k_fast_out int64, config 3 (50k nodes x 100k pods)")
    if o:
        with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
            json.dump({"config": "c3 full", "kernel": "k_fast_out<4,1,0>", "dram_bytes_read": o["dram_read"], "dram_bytes_write": o["dram_write"],
                       "source": "profiles/r2/ncu_full_hot_kernels.json (ncu --set full, one launch)"}, f)
    src = os.path.join(GO, "r2_launches.csv")
    if os.path.exists(src):
        rows = list(csv.reader(open(src)))
        hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
        hdr = rows[hi]
        ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
        with open(os.path.join(OUT, "launches.csv"), "w") as f:
            f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 400 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-latency --no-c4\n")
            w = csv.writer(f)
            w.writerow(["id", "kernel", "ns"])
            for r in rows[hi + 1:]:
                if len(r) > vi:
                    w.writerow([r[0], r[ki].split("(")[0], r[vi].replace(",", "")])
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    main()
