"""Developer helper: run bench.py with the given args and print a condensed line."""
import json
import subprocess
import sys

out = subprocess.run([sys.executable, "bench.py", "--no-cpu-baseline", "--no-latency"] + sys.argv[1:], capture_output=True, text=True)
try:
    d = json.loads(out.stdout.strip().splitlines()[-1])
    print(round(d["ms_per_step"], 3), {k: round(v, 3) for k, v in d["phases_ms"].items()}, "kernel_frac", round(d["roofline"]["frac"], 3),
          "step_frac", round(d["roofline"]["step_frac"], 3), "e2e_ms", round(d["e2e"]["ms_per_step"], 3), d["config"]["path"])
except Exception as e:
    print("bench failed:", e, out.stdout[-2000:], out.stderr[-2000:])
