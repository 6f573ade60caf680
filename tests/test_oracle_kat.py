"""The oracle against the known-answer vectors and against the independent
pure-Python restatement.  CPU only."""
import random

import numpy as np
import pytest
from conftest import kat_nodes, oracle_cluster_from_nodes

from oracle import oracle as ora
from oracle import pyref


def _expect(kat, case):
    names = [n["name"] for n in kat["nodes"]]
    N = len(names)
    feas = [1 if nm in case["feasible"] else 0 for nm in names]
    raw = [case["raw"].get(nm) for nm in names]
    norm = [case["norm"].get(nm, 0) for nm in names]
    chosen = -1 if case["chosen"] is None else names.index(case["chosen"])
    return feas, raw, norm, chosen


def test_static_scores(kat):
    nodes = kat_nodes(kat)
    cl = oracle_cluster_from_nodes(nodes)
    for i, nd in enumerate(kat["nodes"]):
        assert cl.node_static(i) == (nd["alloc_score"], nd["actual_score"])
        assert pyref.allocate_score(nodes[i]) == nd["alloc_score"]
        assert pyref.actual_score(nodes[i]) == nd["actual_score"]


@pytest.mark.parametrize("faithful", [False, True])
def test_kat_cases_c_oracle(kat, faithful):
    cl = oracle_cluster_from_nodes(kat_nodes(kat))
    out = cl.eval([c["labels"] for c in kat["cases"]], faithful=faithful, threads=2)
    for p, case in enumerate(kat["cases"]):
        feas, raw, norm, chosen = _expect(kat, case)
        assert out["feasible"][p].tolist() == feas, case
        assert out["max6"][p].tolist() == case["max6"], case
        for n in range(len(feas)):
            if feas[n]:
                assert out["raw"][p, n] == raw[n], (case, n)
                assert out["norm"][p, n] == norm[n], (case, n)
            else:
                assert out["norm"][p, n] == 0
        assert out["chosen"][p] == chosen
        assert out["count"][p] == sum(feas)


def test_kat_cases_pyref(kat):
    nodes = kat_nodes(kat)
    for case in kat["cases"]:
        feas, raw, norm, chosen = _expect(kat, case)
        r = pyref.cycle(case["labels"], nodes)
        assert [int(f) for f in r["feasible"]] == feas
        assert r["max6"] == case["max6"]
        for n in range(len(feas)):
            if feas[n]:
                assert r["raw"][n] == raw[n] and r["norm"][n] == norm[n]
        assert r["chosen"] == chosen


def test_filter_single_calls(kat):
    cl = oracle_cluster_from_nodes(kat_nodes(kat))
    names = [n["name"] for n in kat["nodes"]]
    for case in kat["cases"]:
        for i, nm in enumerate(names):
            assert cl.filter(i, case["labels"]) == (nm in case["feasible"])


def test_normalize_kat(kat):
    for src, want in kat["normalize"]:
        assert ora.normalize_score(src) == want
        assert pyref.normalize(src)[0] == want
    with pytest.raises(IndexError):
        ora.normalize_score([])


def test_health_string_is_case_sensitive():
    nd = {"has_scv": True, "card_number": 1, "free_sum": 10, "total_sum": 10, "pods": [],
          "cards": [{"health": "healthy", "free": 10, "total": 10, "clock": 1, "core": 1, "bandwidth": 1, "power": 1}]}
    cl = oracle_cluster_from_nodes([nd])
    assert cl.filter(0, {"scv/memory": "1"}) is False     # filter.go:53 compares with "Healthy"
    assert cl.filter(0, {}) is True
    assert pyref.yoda_filter({"scv/memory": "1"}, nd) is False


def test_total_memory_sum_zero_panics():
    nd = {"has_scv": True, "card_number": 1, "free_sum": 0, "total_sum": 0, "pods": [], "cards": []}
    cl = oracle_cluster_from_nodes([nd])
    with pytest.raises(ZeroDivisionError):
        cl.eval([{}])
    with pytest.raises(ZeroDivisionError):
        pyref.cycle({}, [nd])


# ---- random cross-check: C oracle == pure-Python restatement, including adversarial values ----
def _rand_u64(rng, adversarial):
    if not adversarial:
        return rng.choice([0, 1, 2, 500, 1000, 8000, 11000, 40536, rng.randrange(0, 50000)])
    return rng.choice([0, 1, 2, 1000, 2**31, 2**32 - 1, 2**32, 2**62, 2**63 - 1, 2**63, 2**64 - 1, 2**64 // 100, 2**64 // 100 + 1,
                       rng.randrange(0, 2**64), rng.randrange(0, 100000)])


def _rand_label(rng, adversarial):
    pool = ["0", "1", "2", "3", "8", "9", "500", "1000", "8000", "5705", "1215", "7000", "40536"]
    if adversarial:
        pool += ["-1", "-5", "abc", "", "+2", "007", "1_0", "9223372036854775807", "9223372036854775808", "4294967296", "4294967295",
                 "18446744073709551615", " 1"]
    return rng.choice(pool)


def random_nodes(rng, n, adversarial):
    nodes = []
    for i in range(n):
        nc = rng.choice([0, 1, 2, 4, 8, 8, 8, rng.randrange(0, 9)])
        clk = rng.choice([1215, 5705, 7000, 877])
        cards = []
        for g in range(nc):
            cards.append({"health": rng.choice(["Healthy"] * 6 + ["Unhealthy", "healthy"]),
                          "free": _rand_u64(rng, adversarial), "total": _rand_u64(rng, adversarial) or 1,
                          "clock": rng.choice([clk, clk, clk, _rand_u64(rng, adversarial)]),
                          "core": _rand_u64(rng, adversarial), "bandwidth": _rand_u64(rng, adversarial), "power": _rand_u64(rng, adversarial)})
        pods = [rng.choice([{"scv/memory": _rand_label(rng, adversarial)}, {"app": "x"}, {}]) for _ in range(rng.randrange(0, 4))]
        nodes.append({"name": f"n{i}", "has_scv": rng.random() > 0.1, "card_number": rng.choice([nc, nc, 0, 8, _rand_u64(rng, adversarial)]),
                      "cards": cards, "free_sum": _rand_u64(rng, adversarial), "total_sum": _rand_u64(rng, adversarial) or 1, "pods": pods})
    return nodes


def random_pods(rng, n, adversarial):
    out = []
    for _ in range(n):
        d = {}
        if rng.random() < 0.5:
            d["scv/number"] = _rand_label(rng, adversarial)
        if rng.random() < 0.7:
            d["scv/memory"] = _rand_label(rng, adversarial)
        if rng.random() < 0.3:
            d["scv/clock"] = _rand_label(rng, adversarial)
        out.append(d)
    return out


@pytest.mark.parametrize("adversarial", [False, True])
@pytest.mark.parametrize("seed", range(6))
def test_c_oracle_equals_pyref(seed, adversarial):
    rng = random.Random(1000 * seed + adversarial)
    nodes = random_nodes(rng, rng.randrange(1, 24), adversarial)
    pods = random_pods(rng, 12, adversarial)
    cl = oracle_cluster_from_nodes(nodes)
    out = cl.eval(pods, threads=1 + seed % 3)
    for p, lab in enumerate(pods):
        r = pyref.cycle(lab, nodes)
        assert out["feasible"][p].tolist() == [int(f) for f in r["feasible"]]
        assert out["max6"][p].tolist() == r["max6"]
        assert out["raw"][p].tolist() == r["raw"]
        assert out["norm"][p].tolist() == r["norm"]
        assert (out["lowest"][p], out["highest"][p], out["count"][p], out["chosen"][p]) == (r["lowest"], r["highest"], r["count"], r["chosen"])


def test_snapshot_roundtrip_through_oracle(kat):
    """Snapshot (SoA host mirror) -> oracle gives the same answers as the literal AoS construction."""
    from yoda_scheduler_b200.snapshot import Snapshot
    nodes = kat_nodes(kat)
    snap = Snapshot.from_nodes(nodes)
    a = oracle_cluster_from_nodes(nodes).eval([c["labels"] for c in kat["cases"]])
    b = ora.OracleCluster.from_snapshot(snap).eval([c["labels"] for c in kat["cases"]])
    for k in a:
        assert np.array_equal(a[k], b[k]), k
