import time, numpy as np, sys
sys.path.insert(0, '/root/repo')
from yoda_scheduler_b200 import api, synth
snap = synth.make_snapshot(50000)
reqs = synth.make_requests(20000, 5)
with api.Context(50000) as ctx:
    ctx.load_snapshot(snap)
    for P in [int(a) for a in sys.argv[1:]] or (1, 8, 9, 64):
        for flags in (api.EVAL_FULL, api.EVAL_DECIDE):
            r = np.ascontiguousarray(reqs[:P]); ts = []; dev = []
            for i in range(300):
                t0 = time.perf_counter(); ctx.eval(r, flags); ts.append((time.perf_counter() - t0) * 1e3)
                t = ctx.timing(); dev.append((t.total_ms, t.prep_ms, t.max_ms, t.score_ms, t.write_ms, t.h2d_ms, t.d2h_ms, t.kernels))
            d = np.median(np.array(dev[50:]), axis=0)
            print(P, flags, 'wall p50 %.4f' % np.median(ts[50:]), 'dev total %.4f prep %.4f max %.4f score %.4f write %.4f h2d %.4f d2h %.4f kernels %d' % tuple(d))
