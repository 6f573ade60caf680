"""Developer probe: phase times of one shape (device-resident requests, library CUDA events).

  python tools/shape_probe.py c3 | c4 | <nodes>x<pods>  [steps] [raw32]

Run it under `ncu --metrics gpu__time_duration.sum --clock-control none` for a per-kernel launch list of that shape.
"""
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/repo")
from yoda_scheduler_b200 import api, synth  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "c3"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
if shape in synth.CONFIGS:
    n, p, seed = synth.CONFIGS[shape]
else:
    n, p = (int(x) for x in shape.split("x"))
    seed = 3
flags = api.EVAL_FULL | (api.EVAL_RAW32 if "raw32" in sys.argv[3:] else 0)
snap = synth.make_snapshot(n)
reqs = synth.make_requests(p, seed)
d_reqs = torch.from_numpy(reqs.view(np.uint8).copy()).cuda()
d_out = torch.zeros(p * api.POD_RES.itemsize, dtype=torch.uint8, device="cuda")
with api.Context(n) as ctx:
    ctx.load_snapshot(snap)
    rows = []
    for i in range(steps + 3):
        ctx.eval_device(d_reqs.data_ptr(), p, flags, d_out.data_ptr())
        ctx.sync()
        t = ctx.timing()
        rows.append((t.total_ms, t.prep_ms, t.max_ms, t.score_ms, t.write_ms, t.kernels, t.n_classes, t.n_groups))
    med = np.median(np.array(rows[3:]), axis=0)
    print(f"{shape}: {n} nodes x {p} pods  total {med[0]:.3f} prep {med[1]:.3f} max {med[2]:.3f} score {med[3]:.3f} write {med[4]:.3f} ms  "
          f"kernels {int(med[5])} classes {int(med[6])} groups {int(med[7])}")
