"""The C++ host layer: plugin mirror of pkg/yoda/scheduler.go + the fake-framework benchmark harness."""
import json
import subprocess

import numpy as np
import pytest

from yoda_scheduler_b200 import _build, synth
from yoda_scheduler_b200.labels import pods_to_requests

MASK = (1 << 64) - 1


def _fnv(h, v):
    v = int(v)
    for i in range(8):
        h ^= (v >> (8 * i)) & 0xFF
        h = (h * 0x100000001B3) & MASK
    return h


def test_synth_twins_agree():
    """yoda_scheduler_b200/host/synth.hpp and synth.py generate the same cluster and the same pods,
    and the C++ label parsing produces the same requirement vectors as the Python host."""
    N, P, seed = 3000, 500, synth.SEED + 2
    exe = _build.build_host()
    got = json.loads(subprocess.check_output([exe, "--checksum", "--nodes", str(N), "--pods", str(P), "--pod-seed", str(seed)]))
    s = synth.make_snapshot(N)
    h = 0xCBF29CE484222325
    for n in range(N):
        nc = int(s.n_cards[n])
        for v in (s.has_scv[n], s.card_number[n], nc, s.free_memory_sum[n], s.total_memory_sum[n]):
            h = _fnv(h, v)
        for g in range(nc):
            for v in ((int(s.healthy_mask[n]) >> g) & 1, s.free_memory[n, g], s.total_memory[n, g], s.clock[n, g], s.core[n, g], s.bandwidth[n, g], s.power[n, g]):
                h = _fnv(h, v)
        for m in s.pod_mem[s.pod_off[n]:s.pod_off[n + 1]]:
            h = _fnv(h, m)
    assert got["snapshot_digest"] == f"{h:016x}"
    reqs = pods_to_requests(synth.make_pod_labels(P, seed))
    hp = 0xCBF29CE484222325
    for r in reqs:
        for v in (r["flags"], r["number"], r["memory"], r["clock"]):
            hp = _fnv(hp, v)
    assert got["pods_digest"] == f"{hp:016x}"


@pytest.mark.gpu
def test_harness_framework_path_equals_batch_and_oracle(tmp_path):
    """Drive the plugin mirror like kube-scheduler would (Filter per node, PostFilter, Score per feasible
    node, NormalizeScore, selectHost) and in one batch; both must agree with each other (checked inside
    the harness, exit code) and with the oracle (checked here on the dumped per-pod results)."""
    from oracle import oracle as ora
    N, P, seed = 2000, 256, synth.SEED + 2
    exe = _build.build_host()
    dump = tmp_path / "batch.txt"
    r = subprocess.run([exe, "--nodes", str(N), "--pods", str(P), "--pod-seed", str(seed), "--cycles", "24", "--seq", "40", "--dump", str(dump)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    assert info["mismatches"] == 0 and info["cycles_checked"] == 24 and info["path"] == 1
    assert info["seq_pods"] == 40 and info["seq_mismatches"] == 0   # sequential-commit pass == 40 real cycles with assume
    got = np.loadtxt(dump, dtype=np.int64).reshape(P, 5)
    want = ora.OracleCluster.from_snapshot(synth.make_snapshot(N)).eval(synth.make_pod_labels(P, seed), threads=8, full=False)
    assert np.array_equal(got[:, 0], want["chosen"])
    assert np.array_equal(got[:, 1], want["count"])
    assert np.array_equal(got[:, 2], want["lowest"])
    assert np.array_equal(got[:, 3], want["highest"])
    assert np.array_equal(got[:, 4].astype(np.uint64), want["max6"][:, 3])
