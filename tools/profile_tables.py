"""Developer helper: turn an ncu launch list (CSV of gpu__time_duration.sum) and a raw-page CSV of an
`ncu --set full` report into the markdown tables / JSON kept under profiles/.

  python tools/profile_tables.py launches gpurun_out/launches_r1.csv
  ncu -i gpurun_out/prof.ncu-rep --page raw --csv > /tmp/raw.csv
  python tools/profile_tables.py full /tmp/raw.csv profiles/r1_ncu_full_hot_kernels.json
"""
import collections
import csv
import json
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_write.sum", "dram__bytes_read.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__grid_size"]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            agg.setdefault(r[ki].split("(")[0][:60], []).append(float(r[vi].replace(",", "")) / 1e6)   # ns -> ms
        except ValueError:
            pass
    total = sum(sum(v) for v in agg.values())
    print("| kernel | launches | total ms | ms / launch | share |\n|---|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"| `{k}` | {len(v)} | {sum(v):.3f} | {sum(v) / len(v):.4f} | {100 * sum(v) / total:.1f}% |")


def full(path, out_json):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    kn = hdr.index("Kernel Name")
    out = {}
    for r in rows[2:]:
        name = r[kn].split("(")[0]
        out[name] = {w: r[hdr.index(w)] for w in WANT}
        print(f"\n## `{name}`\n\n| metric | value |\n|---|---|")
        for w in WANT:
            print(f"| {w} | {r[hdr.index(w)]} {units[hdr.index(w)]} |")
    json.dump(out, open(out_json, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        full(sys.argv[2], sys.argv[3])
