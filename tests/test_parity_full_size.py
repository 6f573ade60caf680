"""GPU parity at BASELINE.json's full sizes, against the CPU oracle (slow: minutes of oracle time).

config 3: ALL 100 000 pods x 50 000 nodes -- every per-pod output (six maxima, lowest, highest, feasible
count, chosen node) against the oracle, and the three result rows (feasible bits, Score, NormalizeScore)
against the oracle for every pod that sits on a boundary of the hot kernels' walk: first/last pod of each
score-pass chunk and of each output-pass chunk, first pod after a requirement-group change, first pod after
a class change.  Those are the pods a bug in the sweep cache, the mark bits or the chunk bookkeeping
(csrc/fast_kernels.cuh, score_body / out_body) would hit first.

config 4: 1 000 000 nodes, FULL-mode rows of 64 pods against the oracle, from one context and from four
250k-node shards driven through the phase API.
"""
import os

import numpy as np
import pytest

from oracle import oracle as ora
from parity import norm_as_u8
from yoda_scheduler_b200 import api, synth

pytestmark = pytest.mark.gpu

MAX6 = ("max_bandwidth", "max_clock", "max_core", "max_free_memory", "max_power", "max_total_memory")


def _assert_pod_results(res, o, what):
    got_max = np.stack([res[k] for k in MAX6], axis=1)
    assert np.array_equal(got_max, o["max6"]), f"{what}: maxima differ at pods {np.nonzero((got_max != o['max6']).any(axis=1))[0][:8]}"
    for mine, theirs in (("feasible", "count"), ("lowest", "lowest"), ("highest", "highest"), ("chosen", "chosen")):
        bad = np.nonzero(res[mine] != o[theirs])[0]
        assert len(bad) == 0, f"{what}: {mine} differs at {len(bad)} pods, first {bad[:8]}"


def _boundary_pods(orig, grp, cls, score_chunk, out_chunk, cap):
    """Sorted positions on a chunk / group / class boundary -> indices into the caller's batch."""
    P = len(orig)
    pos = set()
    for chunk in (score_chunk, out_chunk):
        for c0 in range(0, P, chunk):
            c1 = min(P, c0 + chunk)
            pos.update((c0, min(c0 + 1, c1 - 1), max(c1 - 2, c0), c1 - 1))
    g_change = np.nonzero(grp[1:] != grp[:-1])[0] + 1
    pos.update(g_change.tolist())
    pos.update((g_change - 1).tolist())
    c_change = np.nonzero(cls[1:] != cls[:-1])[0] + 1
    rng = np.random.default_rng(8)
    if len(c_change) > 400:
        c_change = rng.choice(c_change, 400, replace=False)
    pos.update(c_change.tolist())
    if len(pos) < 2200:     # whatever chunk sizes the runtime picked: never fewer than 2200 pods, the rest drawn at random
        pos.update(rng.choice(P, 2200 - len(pos), replace=False).tolist())
    pos = sorted(pos)
    if len(pos) > cap:
        pos = sorted(rng.choice(pos, cap, replace=False).tolist())
    kinds = {"chunk": len(range(0, P, score_chunk)) + len(range(0, P, out_chunk)), "group": len(g_change), "class": len(c_change)}
    return orig[np.array(pos)], kinds


def test_config_c3_every_pod_and_boundary_rows_vs_oracle():
    n, p, seed = synth.CONFIGS["c3"]
    snap = synth.make_snapshot(n)
    reqs = synth.make_requests(p, seed)
    labels = synth.make_pod_labels(p, seed)
    cl = ora.OracleCluster.from_snapshot(snap)
    threads = os.cpu_count() or 8
    with api.Context(n) as ctx:
        ctx.load_snapshot(snap)
        res = ctx.eval(reqs, api.EVAL_FULL)
        assert ctx.timing().path == 1
        orig, grp, cls, score_chunk, out_chunk = ctx.pod_order(p)
        assert sorted(orig.tolist()) == list(range(p))
        picks, kinds = _boundary_pods(orig, grp, cls, score_chunk, out_chunk, cap=3200)
        assert len(picks) >= 2000, (len(picks), kinds)
        # rows of the boundary pods, against the oracle, 200 pods at a time
        for b0 in range(0, len(picks), 200):
            sel = picks[b0:b0 + 200]
            o = cl.eval([labels[i] for i in sel], threads=threads, full=True)
            want_norm = norm_as_u8(o["norm"])
            for k, i in enumerate(sel):
                assert np.array_equal(ctx.fetch_row(int(i), api.ROW_FEASIBLE_BITS, n), o["feasible"][k]), f"feasible bits of pod {i}"
                assert np.array_equal(ctx.fetch_row(int(i), api.ROW_RAW, n), o["raw"][k]), f"Score row of pod {i}"
                assert np.array_equal(ctx.fetch_row(int(i), api.ROW_NORM, n), want_norm[k]), f"NormalizeScore row of pod {i}"
        dec = ctx.eval(reqs, api.EVAL_DECIDE)      # the one-pass variant must agree with the FULL batch ...
        assert np.array_equal(dec, res)
    # ... and both with the oracle, for every pod of the batch (per-pod outputs; ~2 minutes of host time)
    o = cl.eval(labels, threads=threads, full=False)
    _assert_pod_results(res, o, "config 3, all pods")


def test_config_c4_full_rows_vs_oracle():
    """1M nodes: FULL-mode rows of 64 pods, single context and four shards, against the oracle."""
    import torch
    from yoda_scheduler_b200 import shard
    dev = torch.device("cuda", 0)
    N, P, S = 1_000_000, 64, 4
    snap = synth.make_snapshot(N)
    reqs = synth.make_requests(P, synth.CONFIGS["c4"][2])
    labels = synth.make_pod_labels(P, synth.CONFIGS["c4"][2])
    o = ora.OracleCluster.from_snapshot(snap).eval(labels, threads=os.cpu_count() or 8, full=True)
    want_norm = norm_as_u8(o["norm"])
    with api.Context(N) as single:
        single.load_snapshot(snap)
        res = single.eval(reqs, api.EVAL_FULL)
        assert single.timing().path == 1
        _assert_pod_results(res, o, "config 4, one context")
        for i in range(P):
            assert np.array_equal(single.fetch_row(i, api.ROW_FEASIBLE_BITS, N), o["feasible"][i]), i
            assert np.array_equal(single.fetch_row(i, api.ROW_RAW, N), o["raw"][i]), i
            assert np.array_equal(single.fetch_row(i, api.ROW_NORM, N), want_norm[i]), i
    per = N // S
    ctxs = []
    for r in range(S):
        c = api.Context(per, node_base=r * per, shard_rank=r, shard_count=S)
        c.load_snapshot(snap.slice(r * per, (r + 1) * per))
        c.phase_begin(reqs, api.EVAL_FULL)
        ctxs.append(c)
    for phase in (0, 1, 2):
        for c in ctxs:
            c.phase_run(phase)
        bufs = [c.phase_exchange(phase) for c in ctxs]
        for k in range(len(bufs[0])):
            ts = [shard.device_tensor(b[k][0], b[k][1], dev) for b in bufs]
            red = torch.stack(ts).sum(0) if bufs[0][k][2] == shard.XCHG_SUM else (torch.stack(ts) ^ shard.SIGN).max(0).values ^ shard.SIGN
            for t in ts:
                t.copy_(red)
        torch.cuda.synchronize()
    for r, c in enumerate(ctxs):
        c.phase_run(3)
        _assert_pod_results(c.phase_finish(), o, f"config 4, shard {r}")
        lo, hi = r * per, (r + 1) * per
        for i in range(P):
            assert np.array_equal(c.fetch_row(i, api.ROW_FEASIBLE_BITS, per), o["feasible"][i, lo:hi]), (r, i)
            assert np.array_equal(c.fetch_row(i, api.ROW_RAW, per), o["raw"][i, lo:hi]), (r, i)
            assert np.array_equal(c.fetch_row(i, api.ROW_NORM, per), want_norm[i, lo:hi]), (r, i)
        c.close()
