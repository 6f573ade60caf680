"""Shared helpers of the GPU parity tests: run a batch through the C ABI and through the oracle
and compare every output bit for bit."""
import numpy as np

from oracle import oracle as ora
from yoda_scheduler_b200 import api
from yoda_scheduler_b200.labels import pods_to_requests
from yoda_scheduler_b200.snapshot import Snapshot

I64_MAX = 2**63 - 1


def gpu_cycle(snap: Snapshot, reqs, flags=api.EVAL_FULL, ctx=None, fetch=True):
    own = ctx is None
    if own:
        ctx = api.Context(max(snap.n_nodes, 1))
        ctx.load_snapshot(snap)
    res = ctx.eval(reqs, flags)
    mats = ctx.fetch_all(len(reqs)) if (fetch and flags & api.EVAL_FULL) else None
    t = ctx.timing()
    if own:
        ctx.close()
    return res, mats, t


def norm_as_u8(norm_i64):
    """How the device stores NormalizeScore output: 0..100, or 255 if the int64 rescale wrapped."""
    n = np.asarray(norm_i64)
    return np.where((n >= 0) & (n <= 100), n, 255).astype(np.uint8)


def assert_matches_oracle(snap: Snapshot, pod_labels, res, mats, o=None, what=""):
    if o is None:
        o = ora.OracleCluster.from_snapshot(snap).eval(pod_labels, threads=8)
    P = len(pod_labels)
    got_max = np.stack([res[k] for k in ("max_bandwidth", "max_clock", "max_core", "max_free_memory", "max_power", "max_total_memory")], axis=1)
    assert np.array_equal(got_max, o["max6"]), f"{what} max6 differs at pods {np.nonzero((got_max != o['max6']).any(axis=1))[0][:5]}"
    assert np.array_equal(res["feasible"], o["count"]), f"{what} feasible count"
    assert np.array_equal(res["lowest"], o["lowest"]), f"{what} lowest at {np.nonzero(res['lowest'] != o['lowest'])[0][:5]}"
    assert np.array_equal(res["highest"], o["highest"]), f"{what} highest"
    # node choice: identical whenever the int64 rescale cannot wrap (always, outside adversarial inputs)
    safe = (o["highest"].astype(object) - o["lowest"].astype(object)) <= I64_MAX // 100
    safe = np.array(safe, dtype=bool)
    assert np.array_equal(res["chosen"][safe], o["chosen"][safe]), f"{what} chosen at {np.nonzero(res['chosen'] != o['chosen'])[0][:5]}"
    if mats is not None:
        assert np.array_equal(mats["feasible"], o["feasible"]), f"{what} feasible bits"
        assert np.array_equal(mats["raw"], o["raw"]), f"{what} raw scores"
        assert np.array_equal(mats["norm"], norm_as_u8(o["norm"])), f"{what} normalized scores"
    return o


def labels_and_reqs(pod_labels):
    return pod_labels, pods_to_requests(pod_labels)
