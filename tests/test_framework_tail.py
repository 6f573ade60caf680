"""Framework tail: node sampling, plugin weight, seeded selectHost -- the device version over the
matrices of a FULL batch against the oracle's restatement of kube-scheduler v1.20's generic scheduler."""
import numpy as np
import pytest

from oracle import oracle as ora
from yoda_scheduler_b200 import api, synth
from yoda_scheduler_b200.labels import pods_to_requests
from yoda_scheduler_b200.snapshot import Snapshot

KEYS = ("chosen", "sampled", "processed", "start", "lowest", "highest", "top", "ties")


def test_num_feasible_nodes_to_find_cpu():
    """k8s v1.20 numFeasibleNodesToFind: known values, and the C ABI's host function equals the oracle's."""
    from yoda_scheduler_b200 import _lib
    L = _lib.load()
    assert [ora.num_feasible_nodes_to_find(n) for n in (50, 99, 100, 1000, 5000, 6000, 50000, 1000000)] == [50, 99, 100, 420, 500, 300, 2500, 50000]
    assert ora.num_feasible_nodes_to_find(5000, 30) == 1500 and ora.num_feasible_nodes_to_find(5000, 100) == 5000
    for n in (0, 1, 99, 100, 101, 124, 125, 126, 4999, 5625, 5626, 10**6, 2**31 - 1):
        for pct in (-3, 0, 1, 4, 5, 49, 99, 100, 250):
            assert L.yoda_num_feasible_nodes_to_find(n, pct) == ora.num_feasible_nodes_to_find(n, pct), (n, pct)


def _check(ctx, cl, labels, n_pods, **kw):
    got, nxt = ctx.framework_tail(n_pods, **kw)
    want = cl.framework_tail(labels, **kw)
    for k in KEYS:
        bad = np.nonzero(got[k] != want[k])[0]
        assert len(bad) == 0, f"{k} differs at pods {bad[:6]}: {got[k][bad[:6]]} vs {want[k][bad[:6]]} ({kw})"
    n = ctx.node_count()
    if kw.get("chain"):
        assert nxt == (int(want["start"][-1]) + int(want["processed"][-1])) % n
    return got


@pytest.mark.gpu
@pytest.mark.parametrize("raw32", [False, True])
def test_tail_matches_oracle_on_synthetic_cluster(raw32):
    N, P = 6000, 300     # adaptive percentage: 50 - 48 = 2 -> floor 5% = 300 nodes per cycle
    snap = synth.make_snapshot(N)
    labels = synth.make_pod_labels(P, 77)
    cl = ora.OracleCluster.from_snapshot(snap)
    rng = np.random.default_rng(2)
    other = rng.integers(0, 2000, N).astype(np.int64)
    with api.Context(N) as ctx:
        ctx.load_snapshot(snap)
        ctx.eval(pods_to_requests(labels), api.EVAL_FULL | (api.EVAL_RAW32 if raw32 else 0))
        assert ctx.result_raw_bytes() == (4 if raw32 else 8)
        g = _check(ctx, cl, labels, P, seed=11, chain=True)
        assert (g["sampled"] <= 300).all() and (g["sampled"] == 300).any() and len(np.unique(g["start"])) > 20
        _check(ctx, cl, labels, P, seed=11, chain=False, start_index=4321)
        _check(ctx, cl, labels, P, seed=5, chain=True, start_index=5999, percentage=30, weight=1)
        _check(ctx, cl, labels, P, seed=5, chain=True, percentage=100, other=other)       # every node, other plugins' scores on top
        _check(ctx, cl, labels, P, seed=9, chain=False, percentage=1, other=other, weight=300)


@pytest.mark.gpu
def test_tail_tie_break_is_the_seeded_reservoir():
    """Many identical nodes: most pods end with dozens of nodes on the top total, so the choice is the reservoir's."""
    rng = np.random.default_rng(4)
    nodes = []
    for i in range(1500):
        kind = i % 3
        cards = [{"health": "Healthy", "free": [8000, 8000, 3000][kind], "total": 16000, "clock": 1215, "core": 5000, "bandwidth": 900, "power": 300} for _ in range(4)]
        nodes.append({"name": f"n{i}", "has_scv": i % 50 != 7, "card_number": 4, "cards": cards, "free_sum": sum(c["free"] for c in cards),
                      "total_sum": 64000, "pods": []})
    snap = Snapshot.from_nodes(nodes)
    labels = [{"scv/memory": str(int(m))} for m in rng.integers(1, 9000, 200)] + [{}, {"scv/number": "9"}]
    cl = ora.OracleCluster.from_snapshot(snap)
    with api.Context(1500) as ctx:
        ctx.load_snapshot(snap)
        ctx.eval(pods_to_requests(labels), api.EVAL_FULL)
        for seed in (1, 2, 3):
            g = _check(ctx, cl, labels, len(labels), seed=seed, chain=True, percentage=40)
            assert (g["ties"] > 20).sum() > 100
        a, _ = ctx.framework_tail(len(labels), seed=1, percentage=100)
        b, _ = ctx.framework_tail(len(labels), seed=2, percentage=100)
        assert (a["chosen"] != b["chosen"]).sum() > 50 and np.array_equal(a["ties"], b["ties"])   # the seed moves the choice, nothing else


@pytest.mark.gpu
def test_tail_small_and_degenerate():
    snap = synth.make_snapshot(90)          # < 100 nodes: every node is looked at
    labels = synth.make_pod_labels(40, 3) + [{"scv/memory": "99999999"}]
    cl = ora.OracleCluster.from_snapshot(snap)
    with api.Context(90) as ctx:
        ctx.load_snapshot(snap)
        with pytest.raises(api.YodaError):
            ctx.framework_tail(1)                                  # no FULL batch yet
        ctx.eval(pods_to_requests(labels), api.EVAL_FULL)
        g = _check(ctx, cl, labels, len(labels), seed=3, chain=True)
        assert g["chosen"][-1] == -1 and g["processed"][-1] == 90
        ctx.eval(pods_to_requests(labels), api.EVAL_DECIDE)
        with pytest.raises(api.YodaError):
            ctx.framework_tail(len(labels))                        # DECIDE keeps no matrices
