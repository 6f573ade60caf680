"""The plugin surface (pkg/yoda/scheduler.go) walked the way the scheduling framework walks it:
QueueSort.Less -> Filter per node -> PostFilter -> Score per feasible node -> NormalizeScore,
on the hand-derived known-answer cluster; every answer comes from the device through the C ABI and
is checked against the golden vectors and the pure-Python restatement."""
import pytest
from conftest import kat_nodes

from oracle import pyref

pytestmark = pytest.mark.gpu

def _status_of(node):
    """the Scv.Status document an informer would deliver for a golden node"""
    return {
        "CardNumber": node["card_number"], "FreeMemorySum": node["free_sum"], "TotalMemorySum": node["total_sum"],
        "CardList": [{"Health": c["health"], "FreeMemory": c["free"], "TotalMemory": c["total"], "Clock": c["clock"], "Core": c["core"],
                      "Bandwidth": c["bandwidth"], "Power": c["power"]} for c in node["cards"]],
    }


@pytest.fixture()
def plug(kat):
    from yoda_scheduler_b200 import plugin as pl

    h = pl.Handle()
    y = pl.Yoda(h, max_nodes=64)
    nodes = kat_nodes(kat)
    for nd, spec in zip(nodes, kat["nodes"]):
        name = spec["name"]
        h.node_infos[name] = pl.NodeInfo(name)
        if nd["has_scv"]:
            y.on_scv_upsert(name, _status_of(nd))
        for lb in nd.get("pods", []):
            p = pl.Pod(labels=dict(lb))
            h.node_infos[name].pods.append(p)
            y.on_pod_add(name, p)
    yield pl, y, h, nodes
    y.close()


def test_name(plug):
    pl, y, _, _ = plug
    assert y.name() == "yoda" == pl.NAME


def test_cycle_matches_golden(plug, kat):
    pl, y, h, nodes = plug
    names = [n["name"] for n in kat["nodes"]]
    for case in kat["cases"]:
        pod = pl.Pod(name="p", labels=dict(case["labels"]))
        st = pl.CycleState()
        feasible = []
        for nm in names:                                   # Filter, node by node
            s = y.filter(st, pod, h.node_infos[nm])
            if s.is_success():
                feasible.append(nm)
            else:
                assert s.code == pl.UNSCHEDULABLE and s.message.startswith("Node:" + nm)
        assert feasible == case["feasible"], case
        assert y.post_filter(st, pod).is_success()
        mx = st["Max"]
        assert [mx.MaxBandwidth, mx.MaxClock, mx.MaxCore, mx.MaxFreeMemory, mx.MaxPower, mx.MaxTotalMemory] == case["max6"], case
        if not feasible:
            continue
        scores = []
        for nm in feasible:                                # Score, feasible nodes only
            v, s = y.score(st, pod, nm)
            assert s.is_success() and v == case["raw"][nm], (case, nm)
            scores.append(pl.NodeScore(nm, v))
        assert y.normalize_score(st, pod, scores) is None  # the Go method returns nil
        assert {s.name: s.score for s in scores} == case["norm"], case
        assert max(scores, key=lambda s: s.score).name == case["chosen"]
        assert all(0 <= s.score <= pl.MAX_NODE_SCORE for s in scores)


def test_cycle_matches_pyref(plug):
    pl, y, h, nodes = plug
    for lb in ({"scv/number": "1", "scv/memory": "2000"}, {"scv/clock": "1215"}, {"scv/memory": "-1"}, {"scv/number": "+2", "scv/clock": "7000"}):
        want = pyref.cycle(lb, nodes)
        pod, st = pl.Pod(labels=dict(lb)), pl.CycleState()
        got = [y.filter(st, pod, h.node_infos[n["name"]]).is_success() for n in nodes]
        assert got == [bool(f) for f in want["feasible"]], lb
        y.post_filter(st, pod)
        for i, n in enumerate(nodes):
            if want["feasible"][i]:
                assert y.score(st, pod, n["name"])[0] == want["raw"][i], (lb, n["name"])


def test_filter_without_scv_is_unschedulable(plug):
    pl, y, h, _ = plug
    h.node_infos["ghost"] = pl.NodeInfo("ghost")
    s = y.filter(pl.CycleState(), pl.Pod(labels={}), h.node_infos["ghost"])
    assert s.code == pl.UNSCHEDULABLE                      # scheduler.go:81-84
    assert s.message == 'Node:ghost Scv.core.run-linux.com "ghost" not found'


def test_score_errors(plug):
    pl, y, h, _ = plug
    pod, st = pl.Pod(labels={}), pl.CycleState()
    v, s = y.score(st, pod, "nowhere")                     # scheduler.go:111-114
    assert (v, s.code) == (0, pl.ERROR) and "nodeinfo not found" in s.message
    h.node_infos["ghost"] = pl.NodeInfo("ghost")
    v, s = y.score(st, pod, "ghost")                       # scheduler.go:117-122
    assert (v, s.code) == (0, pl.ERROR) and s.message.startswith("Score Node Error: Scv.core.run-linux.com")
    v, s = y.score(st, pod, "A")                           # algorithm.go:29-33: PostFilter has not written "Max"
    assert (v, s.code) == (0, pl.ERROR) and s.message.startswith("Score Node Error")
    y.post_filter(st, pod)
    v, s = y.score(st, pod, "A")
    assert s.is_success() and v == 891


def test_normalize_foreign_list(plug):
    """The framework's list can be a subset of Yoda's feasible set (another plugin filtered a node out):
    NormalizeScore then rescales exactly the list it is given (scheduler.go:132-157)."""
    pl, y, h, nodes = plug
    pod, st = pl.Pod(labels={}), pl.CycleState()
    for n in nodes:
        y.filter(st, pod, h.node_infos[n["name"]])
    y.post_filter(st, pod)
    lst = [pl.NodeScore(nm, y.score(st, pod, nm)[0]) for nm in ("A", "B")]   # C was dropped by someone else
    y.normalize_score(st, pod, lst)
    assert [s.score for s in lst] == pyref.normalize([891, 1200])[0] == [0, 100]
    one = [pl.NodeScore("B", 1200)]
    y.normalize_score(st, pod, one)
    assert one[0].score == pyref.normalize([1200])[0][0] == 100
    with pytest.raises(IndexError):
        y.normalize_score(st, pod, [])                     # scores[0] on an empty list panics in Go


def test_less(plug):
    pl, y, _, _ = plug
    hi, lo, none, bad = (pl.Pod(labels={"scv/priority": v} if v is not None else {}) for v in ("9", "2", None, "x"))
    assert y.less(hi, lo) and not y.less(lo, hi)           # sort.go:8-10
    assert not y.less(none, bad) and not y.less(bad, none)  # both 0
    assert y.less(lo, none) and not y.less(hi, hi)


def test_informer_events_change_the_next_cycle(plug):
    pl, y, h, nodes = plug
    pod = pl.Pod(labels={"scv/memory": "1000"})

    def raw_of(name):
        st = pl.CycleState()
        for n in nodes:
            y.filter(st, pod, h.node_infos[n["name"]])
        y.post_filter(st, pod)
        return y.score(st, pod, name)[0]

    before = raw_of("B")                                   # 1200 = 700 basic + 300 allocate + 200 actual
    assert before == 1200
    running = pl.Pod(labels={"scv/memory": "11000"})       # half of B's TotalMemorySum -> allocate 150
    y.on_pod_add("B", running)
    assert raw_of("B") == before - 150
    y.on_pod_delete("B", running)
    assert raw_of("B") == before
    y.on_scv_delete("B")
    st = pl.CycleState()
    assert y.filter(st, pod, h.node_infos["B"]).code == pl.UNSCHEDULABLE
    y.on_scv_upsert("B", _status_of(nodes[1]))
    assert raw_of("B") == before


def test_node_registered_mid_cycle_and_refused_scv(plug):
    """A node whose Scv arrives between Filter and Score of a cycle is not part of that cycle's evaluation: Score
    answers "not found" instead of indexing past the memo.  An Scv the device refuses (TotalMemorySum == 0, where
    the reference would panic) takes the node's previous record out of service."""
    pl, y, h, nodes = plug
    from yoda_scheduler_b200 import api
    pod, st = pl.Pod(labels={}), pl.CycleState()
    for n in nodes:
        y.filter(st, pod, h.node_infos[n["name"]])
    y.post_filter(st, pod)
    h.node_infos["late"] = pl.NodeInfo("late")
    y.on_scv_upsert("late", _status_of(nodes[0]))
    v, s = y.score(st, pod, "late")
    assert (v, s.code) == (0, pl.ERROR) and "not found" in s.message
    assert y.filter(st, pod, h.node_infos["late"]).code == pl.UNSCHEDULABLE
    st2 = pl.CycleState()                                  # the next cycle sees it
    assert y.filter(st2, pod, h.node_infos["late"]).is_success()
    broken = _status_of(nodes[0])
    broken["TotalMemorySum"] = 0
    with pytest.raises(api.YodaError):
        y.on_scv_upsert("A", broken)
    st3 = pl.CycleState()
    assert y.filter(st3, pod, h.node_infos["A"]).code == pl.UNSCHEDULABLE
    y.on_scv_upsert("A", _status_of(nodes[0]))
    assert y.filter(pl.CycleState(), pod, h.node_infos["A"]).is_success()
