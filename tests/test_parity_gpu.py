"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle -- bit exact."""
import random

import numpy as np
import pytest
from conftest import kat_nodes

from oracle import oracle as ora
from parity import assert_matches_oracle, gpu_cycle, norm_as_u8
from test_oracle_kat import random_nodes, random_pods
from yoda_scheduler_b200 import api, synth
from yoda_scheduler_b200.labels import pods_to_requests
from yoda_scheduler_b200.snapshot import Snapshot

pytestmark = pytest.mark.gpu

PATHS = [api.EVAL_FULL, api.EVAL_FULL | api.EVAL_FORCE_DIRECT]


@pytest.mark.parametrize("flags", PATHS)
def test_kat_through_c_abi(kat, flags):
    nodes = kat_nodes(kat)
    snap = Snapshot.from_nodes(nodes)
    names = [n["name"] for n in kat["nodes"]]
    labels = [c["labels"] for c in kat["cases"]]
    res, mats, _ = gpu_cycle(snap, pods_to_requests(labels), flags)
    for p, case in enumerate(kat["cases"]):
        feas = [1 if nm in case["feasible"] else 0 for nm in names]
        assert mats["feasible"][p].tolist() == feas
        got_max = [int(res[k][p]) for k in ("max_bandwidth", "max_clock", "max_core", "max_free_memory", "max_power", "max_total_memory")]
        assert got_max == case["max6"]
        for n, nm in enumerate(names):
            if feas[n]:
                assert mats["raw"][p, n] == case["raw"][nm]
                assert mats["norm"][p, n] == case["norm"][nm]
        assert res["chosen"][p] == (-1 if case["chosen"] is None else names.index(case["chosen"]))
    assert_matches_oracle(snap, labels, res, mats)


@pytest.mark.parametrize("flags", PATHS)
@pytest.mark.parametrize("adversarial", [False, True])
@pytest.mark.parametrize("seed", range(24))
def test_random_small(seed, adversarial, flags):
    rng = random.Random(77 * seed + adversarial)
    nodes = random_nodes(rng, rng.choice([1, 3, 31, 32, 33, 127, 129, 300]), adversarial)
    labels = random_pods(rng, rng.choice([1, 2, 17, 129, 400]), adversarial)
    snap = Snapshot.from_nodes(nodes)
    res, mats, _ = gpu_cycle(snap, pods_to_requests(labels), flags)
    assert_matches_oracle(snap, labels, res, mats, what=f"seed={seed} adv={adversarial}")


@pytest.mark.parametrize("flags", PATHS)
def test_config_c1(flags):
    """BASELINE config 1: 100 nodes x 8 cards, one pod scv/memory=1000."""
    snap = synth.make_snapshot(100)
    labels = [{"scv/memory": "1000"}]
    res, mats, _ = gpu_cycle(snap, pods_to_requests(labels), flags)
    assert_matches_oracle(snap, labels, res, mats)
    assert res["chosen"][0] >= 0


@pytest.mark.parametrize("flags", PATHS)
def test_config_c2(flags):
    """BASELINE config 2: 5k nodes, 1k pods with mixed labels."""
    n, p, seed = synth.CONFIGS["c2"]
    snap = synth.make_snapshot(n)
    labels = synth.make_pod_labels(p, seed)
    reqs = synth.make_requests(p, seed)
    assert np.array_equal(reqs, pods_to_requests(labels))
    res, mats, _ = gpu_cycle(snap, reqs, flags)
    assert_matches_oracle(snap, labels, res, mats)


def test_decide_mode_equals_full():
    n, p, seed = synth.CONFIGS["c2"]
    snap = synth.make_snapshot(n)
    reqs = synth.make_requests(p, seed)
    full, _, _ = gpu_cycle(snap, reqs, api.EVAL_FULL, fetch=False)
    dec, _, _ = gpu_cycle(snap, reqs, api.EVAL_DECIDE)
    assert np.array_equal(full, dec)


def test_config_c3_subsample():
    """BASELINE config 3 shape (50k nodes), a 1536-pod slice of the 100k batch, against the oracle."""
    n, p, seed = synth.CONFIGS["c3"]
    snap = synth.make_snapshot(n)
    labels = synth.make_pod_labels(p, seed)[:1536]
    reqs = synth.make_requests(p, seed)[:1536]
    res, mats, _ = gpu_cycle(snap, reqs, api.EVAL_FULL)
    assert_matches_oracle(snap, labels, res, mats)


def test_empty_and_degenerate():
    snap = synth.make_snapshot(64)
    with api.Context(64) as ctx:
        ctx.load_snapshot(snap)
        assert len(ctx.eval(np.zeros(0, api.POD_REQ))) == 0          # empty batch
    # snapshot in which no node has an SCV: nothing is feasible
    s2 = synth.make_snapshot(40)
    s2.has_scv[:] = 0
    labels = [{}, {"scv/memory": "1"}]
    res, mats, _ = gpu_cycle(s2, pods_to_requests(labels))
    assert res["chosen"].tolist() == [-1, -1] and res["feasible"].tolist() == [0, 0]
    assert_matches_oracle(s2, labels, res, mats)
    # a context that never saw a node
    with api.Context(8) as ctx:
        ctx.commit()
        r = ctx.eval(pods_to_requests([{}]), api.EVAL_DECIDE)
        assert r["chosen"][0] == -1 and r["max_core"][0] == 1


@pytest.mark.parametrize("N,P", [(3000, 300), (50000, 384)])
def test_incremental_upserts_equal_full_rebuild(N, P):
    """Churn (BASELINE config 5: 50k nodes, 10% of them updated per tick, interleaved with scheduling
    cycles; also a small shape): the incrementally maintained device snapshot must answer exactly like
    one rebuilt from scratch, and like the oracle."""
    rng = np.random.default_rng(5)
    snap = synth.make_snapshot(N)
    reqs = synth.make_requests(P, synth.SEED + 5)
    labels = synth.make_pod_labels(P, synth.SEED + 5)
    with api.Context(N) as inc:
        inc.load_snapshot(snap)
        for tick in range(3):
            touched = rng.choice(N, N // 10, replace=False)
            other = synth.make_snapshot(N, seed=synth.SEED + 100 + tick)
            for f in ("free_memory", "healthy_mask", "free_memory_sum"):
                getattr(snap, f)[touched] = getattr(other, f)[touched]
            gone = touched[:5]
            snap.has_scv[gone] = 0
            back = touched[5:10]
            snap.has_scv[back] = 1
            alloc = snap.alloc_sum()
            alloc[touched[10:60]] += np.uint64(1024)
            live = touched[snap.has_scv[touched] == 1]
            inc.upsert_many(live, snap.records(live))
            for i in gone:
                inc.remove(int(i))
            inc.set_alloc_many(touched[10:60], alloc[touched[10:60]])
            # eval with staged-but-uncommitted updates is refused
            with pytest.raises(api.YodaError) as e:
                inc.eval(reqs)
            assert e.value.code == -5
            inc.commit()
            r_inc = inc.eval(reqs)
            m_inc = inc.fetch_all(P)
            with api.Context(N) as full:
                full.load_snapshot(snap, commit=False)
                full.set_alloc_many(np.arange(N), alloc)
                full.commit()
                r_full = full.eval(reqs)
                m_full = full.fetch_all(P)
            assert np.array_equal(r_inc, r_full)
            for k in m_inc:
                assert np.array_equal(m_inc[k], m_full[k]), k
            # fold the extra allocations into the host mirror so the oracle sees them too
            extra = np.zeros(N, np.uint64)
            extra[touched[10:60]] = 1024
            counts = np.diff(snap.pod_off).astype(np.int64)
            add = extra > 0
            new_mem, new_off = [], [0]
            for i in range(N):
                seg = list(snap.pod_mem[snap.pod_off[i]:snap.pod_off[i + 1]])
                if add[i]:
                    seg.append(1024)
                new_mem += seg
                new_off.append(len(new_mem))
            snap.pod_mem = np.array(new_mem, np.uint64)
            snap.pod_off = np.array(new_off, np.uint32)
            assert_matches_oracle(snap, labels, r_inc, m_inc, what=f"tick {tick}")


def test_normalize_scores_on_device(kat):
    with api.Context(4) as ctx:
        for src, want in kat["normalize"]:
            assert ctx.normalize_scores(src).tolist() == want
        rng = np.random.default_rng(0)
        for hi in (50, 2**40, 2**62):
            a = rng.integers(-hi, hi, 5000, dtype=np.int64)
            assert ctx.normalize_scores(a).tolist() == ora.normalize_score(a.tolist())
        with pytest.raises(api.YodaError):
            ctx.normalize_scores([])


def test_upsert_validation():
    with api.Context(4) as ctx:
        rec = np.zeros(1, api.NODE_REC)
        rec["total_memory_sum"] = 0
        with pytest.raises(api.YodaError) as e:
            ctx.upsert(0, rec)
        assert e.value.code == -6                      # reference would panic: algorithm.go:71
        rec["total_memory_sum"] = 1
        rec["n_cards"] = 9
        with pytest.raises(api.YodaError) as e:
            ctx.upsert(0, rec)
        assert e.value.code == -3
        rec["n_cards"] = 1
        with pytest.raises(api.YodaError) as e:
            ctx.upsert(4, rec)
        assert e.value.code == -3


def test_staging_ring_overflow_is_safe():
    """More pending upserts than the pinned ring holds: early flushes, still exact after commit."""
    N = 70000  # ring capacity is 65536 records
    snap = synth.make_snapshot(N)
    reqs = synth.make_requests(64, 9)
    labels = synth.make_pod_labels(64, 9)
    res, mats, _ = gpu_cycle(snap, reqs)
    assert_matches_oracle(snap, labels, res, mats)


def test_config_c3_full_size_properties():
    """BASELINE config 3 at full size (50k nodes x 100k pods, 45.6 GB of matrices).  Too big for the oracle
    in a test, so: (1) the fast path must equal the exact-u64 direct kernels on a random 1500-pod subset,
    rows included; (2) size-independent invariants over the whole batch; (3) a second run is identical."""
    n, p, seed = synth.CONFIGS["c3"]
    snap = synth.make_snapshot(n)
    reqs = synth.make_requests(p, seed)
    rng = np.random.default_rng(3)
    pick = np.sort(rng.choice(p, 1500, replace=False))
    with api.Context(n) as ctx:
        ctx.load_snapshot(snap)
        res = ctx.eval(reqs, api.EVAL_FULL)
        assert ctx.timing().path == 1
        rows = {int(i): (ctx.fetch_row(int(i), api.ROW_FEASIBLE_BITS), ctx.fetch_row(int(i), api.ROW_RAW), ctx.fetch_row(int(i), api.ROW_NORM)) for i in pick}
        again = ctx.eval(reqs, api.EVAL_FULL)
        assert np.array_equal(res, again)                                   # idempotent
        sub = ctx.eval(reqs[pick], api.EVAL_FULL | api.EVAL_FORCE_DIRECT)  # exact path, same pods
        assert ctx.timing().path == 0
        assert np.array_equal(sub, res[pick])
        for k, i in enumerate(pick):
            bits, raw, norm = rows[int(i)]
            assert np.array_equal(bits, ctx.fetch_row(k, api.ROW_FEASIBLE_BITS))
            assert np.array_equal(raw, ctx.fetch_row(k, api.ROW_RAW))
            assert np.array_equal(norm, ctx.fetch_row(k, api.ROW_NORM))
            r = res[i]
            assert bits.sum() == r["feasible"]                                # count == popcount of the bit row
            if r["feasible"]:
                f = bits.astype(bool)
                assert raw[f].max() == r["highest"] and norm[f].max() == 100 and norm[~f].max(initial=0) == 0
                assert r["chosen"] == np.nonzero(f & (raw == r["highest"]))[0][0]   # lowest index holding the top score
                assert norm[r["chosen"]] == 100
                lo = r["lowest"]
                want = (raw[f] - lo) * 100 // (r["highest"] - lo)            # scheduler.go:152 (all terms >= 0 here)
                assert np.array_equal(norm[f], want.astype(np.uint8))
            else:
                assert r["chosen"] == -1 and norm.max() == 0
    has = res["feasible"] > 0
    assert (res["chosen"][~has] == -1).all() and (res["chosen"][has] >= 0).all()
    assert (res["highest"] >= res["lowest"]).all()


def test_config_c4_shape_sharded_on_one_gpu():
    """BASELINE config 4 shape: 1M nodes, node-sharded.  Four 250k-node shards (run one after the other on
    this GPU, exchanging partials through the phase API) must equal one 1M-node context and the oracle."""
    import torch
    from yoda_scheduler_b200 import shard
    dev = torch.device("cuda", 0)
    N, P, S = 1_000_000, 384, 4
    snap = synth.make_snapshot(N)
    reqs = synth.make_requests(P, synth.CONFIGS["c4"][2])
    labels = synth.make_pod_labels(P, synth.CONFIGS["c4"][2])
    with api.Context(N) as single:
        single.load_snapshot(snap)
        want = single.eval(reqs, api.EVAL_DECIDE)
    o = ora.OracleCluster.from_snapshot(snap).eval(labels[:96], threads=16, full=False)
    assert np.array_equal(want["chosen"][:96], o["chosen"]) and np.array_equal(want["highest"][:96], o["highest"])
    assert np.array_equal(want["feasible"][:96], o["count"]) and np.array_equal(want["lowest"][:96], o["lowest"])
    per = N // S
    ctxs = []
    for r in range(S):
        c = api.Context(per, node_base=r * per, shard_rank=r, shard_count=S)
        c.load_snapshot(snap.slice(r * per, (r + 1) * per))
        c.phase_begin(reqs, api.EVAL_DECIDE)
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
    for c in ctxs:
        c.phase_run(3)
        assert np.array_equal(c.phase_finish(), want)
        c.close()


def test_fallback_when_groups_exceed_capacity():
    """More than 4096 requirement groups in one batch: the fast path hands the batch to the exact
    direct kernels (decided after grouping).  Results must not care."""
    snap = synth.make_snapshot(257)
    snap.card_number[:] = np.arange(257, dtype=np.uint64) * 40          # spread CardNumber so that number matters
    labels = [{"scv/number": str(i)} for i in range(6000)]               # 6000 distinct groups (number only)
    reqs = pods_to_requests(labels)
    res, mats, t = gpu_cycle(snap, reqs, api.EVAL_FULL)
    assert t.path == 0
    assert_matches_oracle(snap, labels, res, mats)


def test_fallback_when_classes_exceed_capacity():
    """More distinct maxima tuples than class tables: fast maxima phase, then the direct kernels for
    scoring (the switch happens between phases and must be seamless)."""
    n_nodes = 700
    nodes = []
    for i in range(n_nodes):
        cards = [{"health": "Healthy", "free": 8 * i + g + 1, "total": 1_000_000 - (8 * i + g), "clock": 1000, "core": 10, "bandwidth": 10, "power": 10}
                 for g in range(8)]
        nodes.append({"name": f"n{i}", "has_scv": True, "card_number": 8, "cards": cards, "free_sum": sum(c["free"] for c in cards),
                      "total_sum": sum(c["total"] for c in cards), "pods": []})
    snap = Snapshot.from_nodes(nodes)
    labels = [{"scv/memory": str(m)} for m in range(1, 5201)]             # MaxTotalMemory differs for every pod
    reqs = pods_to_requests(labels)
    res, mats, t = gpu_cycle(snap, reqs, api.EVAL_FULL)
    assert t.path == 0 and len(np.unique(res["max_total_memory"])) > 4096
    assert_matches_oracle(snap, labels, res, mats)
    # a batch that fits the tables takes the fast path on the same snapshot
    res2, mats2, t2 = gpu_cycle(snap, reqs[:500], api.EVAL_FULL)
    assert t2.path == 1
    assert_matches_oracle(snap, labels[:500], res2, mats2)


def test_fast_path_is_the_one_that_runs():
    """Guard against a silent fallback: the synthetic configs must use the narrow fast path."""
    snap = synth.make_snapshot(5000)
    reqs = synth.make_requests(1000, synth.SEED + 2)
    with api.Context(5000) as ctx:
        ctx.load_snapshot(snap)
        ctx.eval(reqs, api.EVAL_FULL)
        t = ctx.timing()
        assert t.path == 1 and t.n_groups > 10 and t.n_classes > 5 and t.kernels > 10
        ctx.eval(reqs, api.EVAL_FULL | api.EVAL_FORCE_DIRECT)
        assert ctx.timing().path == 0


@pytest.mark.parametrize("n_pods", [1, 5, 40])
def test_class_table_overflow_falls_back(n_pods):
    """A narrow snapshot whose card scores do not fit the packed class table (Clock*100/MaxBandwidth is
    unbounded, algorithm.go:60): the fast path notices while building the tables -- at the end of the
    evaluation for small batches, before scoring otherwise -- and the exact kernels take over."""
    nodes = [{"name": f"n{i}", "has_scv": True, "card_number": 2, "free_sum": 5000, "total_sum": 16000, "pods": [],
              "cards": [{"health": "Healthy", "free": 4000 + i, "total": 8000, "clock": 30_000_000 + i, "core": 100, "bandwidth": 1 + (i % 3), "power": 50},
                        {"health": "Healthy", "free": 1000, "total": 8000, "clock": 1000, "core": 100, "bandwidth": 2, "power": 50}]}
             for i in range(70)]
    snap = Snapshot.from_nodes(nodes)
    labels = [{"scv/memory": str(500 + 37 * k)} for k in range(n_pods)]
    res, mats, t = gpu_cycle(snap, pods_to_requests(labels), api.EVAL_FULL)
    assert t.path == 0
    assert_matches_oracle(snap, labels, res, mats)


def test_context_reuse_across_batch_sizes_and_modes():
    """One long-lived context, batches of very different sizes and modes back to back (workspace growth,
    small-batch path after the table path and vice versa, sequential mode in between): every answer must
    equal the one a fresh context gives."""
    N = 4000
    snap = synth.make_snapshot(N)
    all_reqs = synth.make_requests(30000, 77)
    plan = [(5, api.EVAL_FULL), (3000, api.EVAL_FULL), (8, api.EVAL_DECIDE), (30000, api.EVAL_DECIDE), (1, api.EVAL_FULL),
            (100, api.EVAL_FULL | api.EVAL_FORCE_DIRECT), (9, api.EVAL_FULL), (2500, api.EVAL_DECIDE), (2, api.EVAL_FULL)]
    with api.Context(N) as ctx:
        ctx.load_snapshot(snap)
        for k, (p, flags) in enumerate(plan):
            reqs = all_reqs[k * 7:k * 7 + p]
            got = ctx.eval(reqs, flags)
            rows = ctx.fetch_all(min(p, 3)) if flags & api.EVAL_FULL else None
            if k == 3:
                seq = ctx.eval_sequential(all_reqs[:500])
            with api.Context(N) as fresh:
                fresh.load_snapshot(snap)
                want = fresh.eval(reqs, flags)
                assert np.array_equal(got, want), (k, p, flags)
                if rows is not None:
                    w = fresh.fetch_all(min(p, 3))
                    for key in rows:
                        assert np.array_equal(rows[key], w[key]), (k, key)
                if k == 3:
                    assert np.array_equal(seq, fresh.eval_sequential(all_reqs[:500]))


@pytest.mark.parametrize("n_pods", [3, 40])
def test_class_tables_plain_division_branches(n_pods):
    """The class tables divide by multiplying with a reciprocal when 100*x fits 32 bits and the maximum
    is at least 2; these nodes report memory in KB (fields above 2^32/100) and every bandwidth is 0 or 1
    (MaxBandwidth == 1, so Clock*100/MaxBandwidth is Clock*100): both plain-division branches run, inside
    the fast path, and must give the reference's quotients."""
    rng = random.Random(5)
    nodes = []
    for i in range(90):
        big = i % 3 != 0                      # every third node stays small: both kinds in one CTA
        total = 3_000_000_000 + 1000 * i if big else 16_000 + i
        cards = [{"health": "Healthy" if rng.random() > 0.1 else "Unhealthy", "free": rng.randrange(total + 1), "total": total,
                  "clock": 1000 + 7 * i + g, "core": 5000 + i, "bandwidth": (i + g) % 2, "power": 300} for g in range(4)]
        nodes.append({"name": f"n{i}", "has_scv": True, "card_number": 4, "free_sum": sum(c["free"] for c in cards),
                      "total_sum": 4 * total, "pods": [], "cards": cards})
    snap = Snapshot.from_nodes(nodes)
    labels = [{"scv/memory": str(rng.choice([1, 9000, 40_000_000, 1_500_000_000, 2_900_000_000])), **({"scv/number": "2"} if k % 2 else {})}
              for k in range(n_pods)]
    labels[0] = {"scv/memory": "4000000000"}  # nothing fits: the all-ones class
    res, mats, t = gpu_cycle(snap, pods_to_requests(labels), api.EVAL_FULL)
    assert t.path == 1, "these values are narrow: the table path must have handled them"
    assert_matches_oracle(snap, labels, res, mats)


def test_class_tables_families_and_plain_classes_interleaved():
    """The class tables are computed family by family: classes sorted by their maxima, quotients and admission
    bits carried from one class to the next and recomputed only for the maxima that change.  Four card models
    make families that differ in every divisor; inside a family the classes differ in MaxFreeMemory only, in
    MaxClock only (the D nodes), or in both; and the pods that ask for clock 1200 see only the B cards, whose
    power is 1 -- a plain-division class that sorts into the middle of the walk, between reciprocal-path classes."""
    rng = random.Random(77)
    models = {"A": dict(bandwidth=100, clock=800, core=1, power=200, total=10000),
              "B": dict(bandwidth=500, clock=1200, core=4000, power=1, total=20000),
              "C": dict(bandwidth=900, clock=1000, core=6000, power=300, total=40000),
              "D": dict(bandwidth=300, clock=900, core=5000, power=250, total=80000)}
    nodes = []
    for i in range(240):
        kind = "ABCD"[i % 4]
        m = dict(models[kind])
        if kind == "D":
            m["clock"] = 900 + 10 * (i % 7)     # MaxClock moves with the feasible set while the other maxima stay
        nc = rng.choice([2, 4, 8])
        cards = [{"health": "Healthy" if rng.random() > 0.1 else "Unhealthy", "free": rng.randrange(m["total"] + 1), **m} for _ in range(nc)]
        nodes.append({"name": f"n{i}", "has_scv": True, "card_number": nc, "cards": cards, "free_sum": sum(c["free"] for c in cards),
                      "total_sum": nc * m["total"], "pods": []})
    labels = []
    for k in range(400):
        top = rng.choice([10000, 20000, 40000, 80000])
        d = {"scv/memory": str(rng.randrange(top // 2, top)), "scv/number": str(rng.choice([1, 1, 2, 3, 4]))}
        if k % 5 == 0:
            d["scv/clock"] = str(rng.choice([1200, 1000, 940, 920]))
        labels.append(d)
    snap = Snapshot.from_nodes(nodes)
    res, mats, t = gpu_cycle(snap, pods_to_requests(labels), api.EVAL_FULL)
    assert t.path == 1 and t.n_classes >= 15, (t.path, t.n_classes)
    o = assert_matches_oracle(snap, labels, res, mats, what="class families")
    maxima = np.unique(np.stack([res[k] for k in ("max_bandwidth", "max_clock", "max_core", "max_free_memory", "max_power", "max_total_memory")], axis=1), axis=0)
    plain = maxima[(maxima <= 1).any(axis=1) & (maxima[:, 0] > 1)]
    assert len(plain) >= 2, "expected plain-division classes with a bandwidth maximum above 1 (they sort between the families)"


def test_maxima_scan_across_tiles():
    """Requirement groups much longer than one scan tile (1024 pods), groups that start and end inside a
    tile, and a tail of one-pod groups: the maxima of every pod against the oracle."""
    rng = random.Random(11)
    N, P = 300, 5000
    snap = synth.make_snapshot(N)
    labels = []
    for k in range(P):
        r = rng.random()
        if r < 0.70:
            labels.append({"scv/memory": str(rng.randrange(1, 90000))})                       # one group of ~3500 pods
        elif r < 0.90:
            labels.append({"scv/memory": str(rng.randrange(1, 90000)), "scv/number": "2"})    # ~1000
        elif r < 0.98:
            labels.append({"scv/number": str(rng.choice([1, 4]))})                            # no memory: all equal inside the group
        else:
            labels.append({"scv/clock": str(1000 + k), "scv/memory": str(rng.randrange(1, 90000))})  # ~100 groups of one pod
    res, mats, t = gpu_cycle(snap, pods_to_requests(labels), api.EVAL_DECIDE, fetch=False)
    assert t.path == 1 and t.n_groups > 50
    assert_matches_oracle(snap, labels, res, None)


def test_raw32_matrix_equals_int64_matrix():
    """YODA_EVAL_RAW32: the Score matrix stored as uint32.  Same per-pod results, same rows through
    yoda_fetch_row, and read straight from the device the uint32 matrix is the int64 one narrowed."""
    import torch
    from yoda_scheduler_b200 import shard
    N, P = 3000, 700
    snap = synth.make_snapshot(N)
    reqs = synth.make_requests(P, 5)
    with api.Context(N) as ctx:
        ctx.load_snapshot(snap)
        want = ctx.eval(reqs, api.EVAL_FULL)
        assert ctx.result_raw_bytes() == 8
        want_m = ctx.fetch_all(P)
        got = ctx.eval(reqs, api.EVAL_FULL | api.EVAL_RAW32)
        assert ctx.result_raw_bytes() == 4 and ctx.timing().path == 1
        got_m = ctx.fetch_all(P)
        assert np.array_equal(got, want)
        for k in want_m:
            assert np.array_equal(got_m[k], want_m[k]), k
        import ctypes as C
        d_raw, stride = C.c_void_p(), C.c_uint64()
        api.check(ctx.L.yoda_result_matrices(ctx.h, None, C.byref(d_raw), None, C.byref(stride)))
        t = shard.device_tensor(d_raw.value, P * stride.value // 2, torch.device("cuda", 0))   # uint64 words = 2 scores each
        dev = t.cpu().numpy().view(np.uint32).reshape(P, stride.value)[:, :N]
        assert np.array_equal(dev.astype(np.int64), want_m["raw"])
        # a forced-direct batch cannot honour the request: int64 rows, same values
        got = ctx.eval(reqs[:50], api.EVAL_FULL | api.EVAL_RAW32 | api.EVAL_FORCE_DIRECT)
        assert ctx.result_raw_bytes() == 8
        assert np.array_equal(got, want[:50])
        assert np.array_equal(ctx.fetch_all(50)["raw"], want_m["raw"][:50])
    assert_matches_oracle(snap, synth.make_pod_labels(P, 5), want, want_m)


def test_raw32_request_survives_a_fallback():
    """The class-table overflow case with YODA_EVAL_RAW32: the exact kernels take over, the rows are int64."""
    nodes = [{"name": f"n{i}", "has_scv": True, "card_number": 2, "free_sum": 5000, "total_sum": 16000, "pods": [],
              "cards": [{"health": "Healthy", "free": 4000 + i, "total": 8000, "clock": 30_000_000 + i, "core": 100, "bandwidth": 1 + (i % 3), "power": 50},
                        {"health": "Healthy", "free": 1000, "total": 8000, "clock": 1000, "core": 100, "bandwidth": 2, "power": 50}]}
             for i in range(70)]
    snap = Snapshot.from_nodes(nodes)
    for n_pods in (3, 40):
        labels = [{"scv/memory": str(500 + 37 * k)} for k in range(n_pods)]
        with api.Context(70) as ctx:
            ctx.load_snapshot(snap)
            res = ctx.eval(pods_to_requests(labels), api.EVAL_FULL | api.EVAL_RAW32)
            assert ctx.timing().path == 0 and ctx.result_raw_bytes() == 8
            assert_matches_oracle(snap, labels, res, ctx.fetch_all(n_pods))
    with api.Context(70) as ctx:
        ctx.load_snapshot(snap)
        with pytest.raises(api.YodaError):
            ctx.eval(pods_to_requests(labels), 64)          # unknown flag bits are rejected


@pytest.mark.parametrize("seed", [1, 2])
def test_heterogeneous_cards_medium(seed):
    """Nodes whose cards differ in every field (so several cards per node survive the dominance pruning of
    the maxima pass, and nodes of one warp disagree on clock tests), many requirement groups, on the fast
    path: every output against the oracle."""
    rng = random.Random(900 + seed)
    N, P = 1500, 2500
    clocks = [877, 1215, 1410, 5705, 7000]
    nodes = []
    for i in range(N):
        nc = rng.choice([1, 2, 4, 8, 8])
        cards = []
        for g in range(nc):
            total = rng.choice([8000, 11000, 16000, 24000, 40536, 81920])
            cards.append({"health": "Healthy" if rng.random() > 0.08 else "Unhealthy", "free": rng.randrange(total + 1), "total": total,
                          "clock": rng.choice(clocks), "core": rng.choice([2560, 3584, 4352, 5120, 6912]),
                          "bandwidth": rng.choice([320, 352, 448, 616, 900, 1555]), "power": rng.choice([70, 250, 260, 300, 400])})
        running = [{"scv/memory": str(rng.randrange(100, 30000))} for _ in range(rng.randrange(0, 3))]
        nodes.append({"name": f"n{i}", "has_scv": rng.random() > 0.02, "card_number": rng.choice([nc, nc, nc, 8, 0]), "cards": cards,
                      "free_sum": sum(c["free"] for c in cards), "total_sum": sum(c["total"] for c in cards), "pods": running})
    labels = []
    for _ in range(P):
        d = {}
        if rng.random() < 0.6:
            d["scv/number"] = str(rng.choice([0, 1, 1, 2, 2, 3, 4, 8, 9]))
        if rng.random() < 0.8:
            d["scv/memory"] = str(rng.randrange(1, 82000))
        if rng.random() < 0.4:
            d["scv/clock"] = str(rng.choice(clocks + [999]))
        labels.append(d)
    snap = Snapshot.from_nodes(nodes)
    res, mats, t = gpu_cycle(snap, pods_to_requests(labels), api.EVAL_FULL)
    assert t.path == 1 and t.n_groups > 60
    assert_matches_oracle(snap, labels, res, mats, what=f"heterogeneous seed={seed}")


def test_cluster_grows_and_cards_are_replaced():
    """Between two cycles the cluster gains nodes and some nodes get different cards (other clocks, cores,
    memory sizes): everything derived at commit -- narrow planes, node order by clock, surviving cards,
    k-th largest FreeMemory -- has to follow.  Checked against a fresh context and the oracle."""
    N0, N1, P = 1000, 1500, 600
    final = synth.make_snapshot(N1)
    other = synth.make_snapshot(N1, seed=synth.SEED + 9)     # same shapes, different SKUs per node
    reqs = synth.make_requests(P, 41)
    labels = synth.make_pod_labels(P, 41)
    with api.Context(N1) as ctx:
        first = np.arange(N0)
        live = first[other.has_scv[first] == 1]
        ctx.upsert_many(live, other.records(live))            # cycle 1 sees the "other" cards on nodes 0..N0-1
        ctx.commit()
        assert ctx.node_count() == int(live.max()) + 1
        ctx.eval(reqs, api.EVAL_FULL)
        assert ctx.timing().path == 1
        ctx.load_snapshot(final)                              # replace every node's cards, add nodes N0..N1-1
        got = ctx.eval(reqs, api.EVAL_FULL)
        got_m = ctx.fetch_all(P)
        assert ctx.timing().path == 1
    with api.Context(N1) as fresh:
        fresh.load_snapshot(final)
        want = fresh.eval(reqs, api.EVAL_FULL)
        want_m = fresh.fetch_all(P)
    assert got_m["raw"].shape == want_m["raw"].shape
    assert np.array_equal(got, want)
    for k in got_m:
        assert np.array_equal(got_m[k], want_m[k]), k
    assert_matches_oracle(final, labels, got, got_m, what="grown cluster")


def _hetero_nodes(rng, n):
    """Nodes whose cards differ field by field and whose FreeMemory values collide across nodes (few distinct
    values), so that lanes of one warp land on the same or on neighbouring pod positions in the maxima searches."""
    nodes = []
    for i in range(n):
        nc = rng.choice([1, 2, 3, 8])
        cards = [{"health": "Healthy" if rng.random() > 0.15 else "Unhealthy", "free": rng.choice([0, 100, 100, 2000, 2000, 7777, 16000, 40000]),
                  "total": rng.choice([16000, 40536]), "clock": rng.choice([877, 1215, 1215, 5705]), "core": rng.randrange(1, 7000),
                  "bandwidth": rng.randrange(0, 2000), "power": rng.randrange(1, 500)} for _ in range(nc)]
        nodes.append({"name": f"n{i}", "has_scv": rng.random() > 0.05, "card_number": rng.choice([nc, 8, 0]), "cards": cards,
                      "free_sum": sum(c["free"] for c in cards), "total_sum": sum(c["total"] for c in cards), "pods": []})
    return nodes


@pytest.mark.parametrize("n_nodes", [31, 33, 257])
def test_maxima_lane_teaming_stress(n_nodes):
    """k_max_thresholds lets whichever lanes of a warp arrive together at the same pod position share one atomic
    (teams formed from __activemask() inside divergent code).  Nothing but results can catch a mis-teamed
    reduction, so: tiny groups (1-3 pods, plus 9-12 pods to leave the small-batch path), heterogeneous nodes with
    colliding FreeMemory values, 200 seeds per shape, every maximum against the oracle."""
    with api.Context(n_nodes) as ctx:
        for seed in range(200):
            rng = random.Random(n_nodes * 1000 + seed)
            snap = Snapshot.from_nodes(_hetero_nodes(rng, n_nodes))
            ctx.load_snapshot(snap)
            n_pods = rng.choice([1, 2, 3, 3, 9, 12])
            labels = []
            for _ in range(n_pods):
                d = {"scv/memory": str(rng.choice([0, 1, 100, 101, 2000, 2001, 7777, 16000, 39999, 40001]))} if rng.random() < 0.8 else {}
                if rng.random() < 0.4:
                    d["scv/number"] = str(rng.choice([1, 2, 3]))
                if rng.random() < 0.3:
                    d["scv/clock"] = str(rng.choice([877, 1215, 5705]))
                labels.append(d)
            res = ctx.eval(pods_to_requests(labels), api.EVAL_DECIDE)
            assert ctx.timing().path == 1
            assert_matches_oracle(snap, labels, res, None, what=f"nodes={n_nodes} seed={seed}")


def _mixed_snapshot(N, kinds, rng):
    """A synthetic snapshot in which a few nodes carry a value outside the narrow (uint32) range."""
    snap = synth.make_snapshot(N)
    have = np.nonzero((snap.has_scv == 1) & (snap.n_cards > 0))[0]
    picks = rng.choice(have, len(kinds), replace=False)
    for n, kind in zip(picks, kinds):
        if kind == "cardnum":      # Status.CardNumber is a Go uint: any 64-bit value is legal (filter.go:11-16)
            snap.card_number[n] = np.uint64(1) << np.uint64(40)
        elif kind == "stat":       # FreeMemorySum far above TotalMemorySum: the static score leaves 24 bits (algorithm.go:71)
            snap.free_memory_sum[n] = np.uint64(1) << np.uint64(50)
        elif kind == "stat_huge":  # ... and 63 bits: uint64 -> int64 turns the score into 0 (filter.go:84-86)
            snap.free_memory_sum[n] = np.uint64(2**64 - 1) // np.uint64(100)
            snap.total_memory_sum[n] = np.uint64(1)
        elif kind == "power":      # a card field beyond 32 bits that is not a divisor of anyone else's score... but it is
            snap.power[n, 0] = np.uint64(1) << np.uint64(33)   # a maximum: the classes of the pods it fits go wide
        elif kind == "free":
            snap.free_memory[n, 0] = np.uint64(1) << np.uint64(36)
    return snap, picks


@pytest.mark.parametrize("flags", [api.EVAL_FULL, api.EVAL_DECIDE])
@pytest.mark.parametrize("kinds", [("cardnum",), ("stat",), ("stat_huge",), ("cardnum", "stat", "stat", "cardnum"), ("power",), ("free", "cardnum")])
def test_mixed_wide_and_narrow_nodes(kinds, flags):
    """One misreported SCV field must not cost the whole batch the table path: wide nodes are evaluated by the
    exact kernels, every other node by the tables, in the same batch, and the result is the oracle's.  Wide card
    fields that become a class maximum (power, free) still send the batch to the exact kernels as a whole."""
    rng = np.random.default_rng(len(kinds) * 7 + flags)
    N, P = 2100, 700
    snap, picks = _mixed_snapshot(N, kinds, rng)
    labels = synth.make_pod_labels(P, 19)
    reqs = pods_to_requests(labels)
    res, mats, t = gpu_cycle(snap, reqs, flags)
    wide_maxima = any(k in ("power", "free") for k in kinds)
    assert t.path == (0 if wide_maxima else 1), "wide nodes alone must not force the exact kernels on everyone"
    o = assert_matches_oracle(snap, labels, res, mats, what=f"mixed {kinds}")
    if "stat" in kinds and flags & api.EVAL_FULL:
        # the wide node's score is the highest of every pod it fits: the 32-bit rescale cannot hold, the rows still match
        assert (o["highest"] >= 2**32).any()
    # small batches (their own code path) and RAW32 (a request the mixed snapshot cannot honour: int64 rows)
    with api.Context(N) as ctx:
        ctx.load_snapshot(snap)
        r5 = ctx.eval(reqs[:5], flags)
        assert np.array_equal(r5, res[:5])
        if flags & api.EVAL_FULL:
            m5 = ctx.fetch_all(5)
            for k in m5:
                assert np.array_equal(m5[k], mats[k][:5]), k
            r32 = ctx.eval(reqs, flags | api.EVAL_RAW32)
            assert ctx.result_raw_bytes() == 8 and np.array_equal(r32, res)
            assert np.array_equal(ctx.fetch_row(3, api.ROW_RAW, N), mats["raw"][3])
        # the node is repaired: the snapshot is narrow again and the wide list empties
        fixed = synth.make_snapshot(N)
        live = picks[fixed.has_scv[picks] == 1]
        ctx.upsert_many(live, fixed.records(live))
        ctx.commit()
        again = ctx.eval(reqs, flags | api.EVAL_RAW32)
        assert ctx.timing().path == 1
        if flags & api.EVAL_FULL:
            assert ctx.result_raw_bytes() == 4
        want_fixed = ora.OracleCluster.from_snapshot(fixed).eval(labels, threads=8, full=False)
        assert np.array_equal(again["chosen"], want_fixed["chosen"]) and np.array_equal(again["highest"], want_fixed["highest"])
        assert np.array_equal(again["feasible"], want_fixed["count"])


def test_mostly_wide_snapshot_goes_direct():
    snap = synth.make_snapshot(600)
    snap.card_number[:] = np.uint64(1) << np.uint64(40)
    labels = synth.make_pod_labels(50, 3)
    res, mats, t = gpu_cycle(snap, pods_to_requests(labels), api.EVAL_FULL)
    assert t.path == 0
    assert_matches_oracle(snap, labels, res, mats)


def test_counts_left_on_the_device_and_their_redo():
    """From its second batch on a context sizes nothing from read-backs: group and class counts stay on the device and a
    batch that does not fit what earlier batches sized (more classes than the table buffer holds, more classes than any
    buffer may hold, more groups than the limit) is flagged there and redone with the counts read back.  Every batch of
    the sequence against the oracle, on one long-lived context."""
    n_nodes = 700
    nodes = []
    for i in range(n_nodes):
        cards = [{"health": "Healthy", "free": 8 * i + g + 1, "total": 1_000_000 - (8 * i + g), "clock": 1000, "core": 10, "bandwidth": 10, "power": 10}
                 for g in range(8)]
        nodes.append({"name": f"n{i}", "has_scv": True, "card_number": 8 + 40 * i, "cards": cards, "free_sum": sum(c["free"] for c in cards),
                      "total_sum": sum(c["total"] for c in cards), "pods": []})
    snap = Snapshot.from_nodes(nodes)
    many_classes = [{"scv/memory": str(m)} for m in range(1, 5201)]      # MaxTotalMemory differs for every pod
    many_groups = [{"scv/number": str(i)} for i in range(6000)]          # one requirement group per pod
    plan = [("few classes", many_classes[:30], 1), ("more classes than the buffer", many_classes[:500], 1), ("again, now it fits", many_classes[100:600], 1),
            ("more classes than any buffer", many_classes, 0), ("back to few", many_classes[:40], 1), ("too many groups", many_groups, 0),
            ("and a normal batch", many_classes[:300], 1)]
    with api.Context(n_nodes) as ctx:
        ctx.load_snapshot(snap)
        for what, labels, path in plan:
            for flags in (api.EVAL_FULL, api.EVAL_DECIDE):
                res = ctx.eval(pods_to_requests(labels), flags)
                assert ctx.timing().path == path, what
                mats = ctx.fetch_all(min(len(labels), 60)) if flags & api.EVAL_FULL else None
                o = ora.OracleCluster.from_snapshot(snap).eval(labels, threads=8)
                if mats is not None:
                    o_small = {k: (v[:60] if k in ("feasible", "raw", "norm") else v) for k, v in o.items()}
                    assert_matches_oracle(snap, labels, res, mats, o=o_small, what=what)
                else:
                    assert_matches_oracle(snap, labels, res, None, o=o, what=what)


def test_eval_without_reading_results_back():
    """yoda_eval_batch(out = NULL): the batch is evaluated and its matrices kept; nothing is copied to the host.  What a node
    shard does when another shard's host consumes the (identical) per-pod results."""
    snap = synth.make_snapshot(900)
    reqs = synth.make_requests(70, 13)
    with api.Context(900) as ctx:
        ctx.load_snapshot(snap)
        want = ctx.eval(reqs, api.EVAL_FULL)
        want_rows = ctx.fetch_all(5)
        for n in (70, 3):
            api.check(ctx.L.yoda_eval_batch(ctx.h, reqs.ctypes.data, n, api.EVAL_FULL, None))
            rows = ctx.fetch_all(min(n, 5))
            for k in rows:
                assert np.array_equal(rows[k], want_rows[k][:min(n, 5)]), k
        assert np.array_equal(ctx.eval(reqs, api.EVAL_FULL), want)


def test_single_pod_cycles_replayed_from_a_graph():
    """The one-pod scheduling cycle is captured into a CUDA graph on its second occurrence and replayed afterwards.  The
    replays must follow the data: different pods, a commit that changes the metrics (same node count), a node going wide,
    and a snapshot whose class table overflows (the verdict read back after the replay sends the cycle to the exact
    kernels).  Every cycle against the oracle, rows included."""
    rng = np.random.default_rng(17)
    N = 1200
    snap = synth.make_snapshot(N)
    labels = synth.make_pod_labels(40, 23)
    with api.Context(N) as ctx:
        ctx.load_snapshot(snap)
        cl = ora.OracleCluster.from_snapshot(snap)

        def cycle(lab, flags=api.EVAL_FULL, path=1):
            res = ctx.eval(pods_to_requests(lab), flags)
            assert ctx.timing().path == path
            mats = ctx.fetch_all(len(lab)) if flags & api.EVAL_FULL else None
            assert_matches_oracle(snap, lab, res, mats, o=cl.eval(lab, threads=2))
        for k in range(8):                                   # call by call, capture, then replays
            cycle(labels[k:k + 1])
        for k in range(4):                                   # another key (3 pods), again captured on its second cycle
            cycle(labels[10 + k:13 + k])
        for k in range(4):
            cycle(labels[20 + k:21 + k], api.EVAL_DECIDE)
        # a metric tick: same nodes, new FreeMemory -> the cached graph must read the new snapshot
        other = synth.make_snapshot(N, seed=synth.SEED + 4)
        touched = rng.choice(N, N // 5, replace=False)
        for f in ("free_memory", "healthy_mask", "free_memory_sum"):
            getattr(snap, f)[touched] = getattr(other, f)[touched]
        live = touched[snap.has_scv[touched] == 1]
        ctx.upsert_many(live, snap.records(live))
        ctx.commit()
        cl = ora.OracleCluster.from_snapshot(snap)
        for k in range(4):
            cycle(labels[30 + k:31 + k])
        # a big FULL batch in between reallocates pod buffers, matrices and class tables: the cached graph must not survive it
        big = synth.make_pod_labels(6000, 29)
        res = ctx.eval(pods_to_requests(big), api.EVAL_FULL)
        assert_matches_oracle(snap, big, res, None, o=cl.eval(big, threads=8, full=False))
        for k in range(4):
            cycle(labels[k:k + 1])
        # a node goes wide: another key (wide-node kernels join the sequence)
        node = int(np.nonzero((snap.has_scv == 1) & (snap.n_cards == 8))[0][2])
        snap.card_number[node] = np.uint64(1) << np.uint64(40)
        ctx.upsert_many(np.array([node], np.uint32), snap.records(np.array([node])))
        ctx.commit()
        cl = ora.OracleCluster.from_snapshot(snap)
        for k in range(4):
            cycle(labels[k:k + 1])
    # class-table overflow: every cycle ends on the exact kernels, replayed or not
    nodes = [{"name": f"n{i}", "has_scv": True, "card_number": 2, "free_sum": 5000, "total_sum": 16000, "pods": [],
              "cards": [{"health": "Healthy", "free": 4000 + i, "total": 8000, "clock": 30_000_000 + i, "core": 100, "bandwidth": 1 + (i % 3), "power": 50},
                        {"health": "Healthy", "free": 1000, "total": 8000, "clock": 1000, "core": 100, "bandwidth": 2, "power": 50}]}
             for i in range(70)]
    snap2 = Snapshot.from_nodes(nodes)
    with api.Context(70) as ctx:
        ctx.load_snapshot(snap2)
        for k in range(5):
            lab = [{"scv/memory": str(500 + 37 * k)}]
            res = ctx.eval(pods_to_requests(lab), api.EVAL_FULL)
            assert ctx.timing().path == 0
            assert_matches_oracle(snap2, lab, res, ctx.fetch_all(1))
