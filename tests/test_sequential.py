"""Sequential-commit batch semantics: a run of single-pod cycles in which every
chosen node's allocateMemorySum grows before the next pod is scored (algorithm.go:74-87)."""
import numpy as np
import pytest

from oracle import oracle as ora
from yoda_scheduler_b200 import synth
from yoda_scheduler_b200.labels import pods_to_requests


def _cluster(n):
    snap = synth.make_snapshot(n)
    return snap, ora.OracleCluster.from_snapshot(snap)


def test_oracle_sequential_differs_from_frozen_and_restores():
    snap, cl = _cluster(200)
    labels = [{"scv/memory": "20000"} for _ in range(40)]            # identical pods: frozen = same node 40 times
    frozen = cl.eval(labels, full=False)
    seq = cl.eval_sequential(labels)
    assert len(set(frozen["chosen"].tolist())) == 1
    assert len(set(seq["chosen"].tolist())) > 1                       # allocations push later pods elsewhere
    assert seq["chosen"][0] == frozen["chosen"][0]
    assert np.array_equal(cl.eval(labels, full=False)["chosen"], frozen["chosen"])   # cluster restored
    # pods without scv/memory commit nothing
    plain = [{"scv/number": "2"}] * 10
    assert np.array_equal(cl.eval_sequential(plain)["chosen"], cl.eval(plain, full=False)["chosen"])


@pytest.mark.gpu
@pytest.mark.parametrize("n_nodes,n_pods", [(64, 30), (700, 400), (5000, 600)])
def test_device_sequential_matches_oracle(n_nodes, n_pods):
    from yoda_scheduler_b200 import api
    snap, cl = _cluster(n_nodes)
    labels = synth.make_pod_labels(n_pods, 123)
    for i in range(0, n_pods, 3):                                      # make allocations matter: big memory labels
        labels[i] = dict(labels[i], **{"scv/memory": str(3000 + 7 * i)})
    reqs = pods_to_requests(labels)
    rng = np.random.default_rng(1)
    for order in (None, rng.permutation(n_pods).astype(np.uint32)):
        want = cl.eval_sequential(labels, order)
        with api.Context(n_nodes) as ctx:
            ctx.load_snapshot(snap)
            got = ctx.eval_sequential(reqs, order)
            assert np.array_equal(got["chosen"], want["chosen"])
            assert np.array_equal(got["highest"], want["highest"])
            assert np.array_equal(got["lowest"], want["lowest"])
            assert np.array_equal(got["feasible"], want["count"])
            assert np.array_equal(got["max_free_memory"], want["max6"][:, 3])
            # the committed snapshot was not touched: a frozen batch afterwards equals the oracle's frozen batch
            frozen = ctx.eval(reqs, api.EVAL_DECIDE)
            assert np.array_equal(frozen["chosen"], cl.eval(labels, full=False)["chosen"])
    assert len(set(want["chosen"].tolist())) > 5
