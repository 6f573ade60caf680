"""Developer soak: many more seeds of the small random parity cases than the test suite runs (GPU vs oracle, FULL mode,
table path and exact path), plus medium heterogeneous clusters.  usage: python tools/soak_parity.py [first_seed] [n_seeds]"""
import random
import sys

sys.path.insert(0, "/root/repo")
sys.path.insert(0, "/root/repo/tests")
from parity import assert_matches_oracle, gpu_cycle  # noqa: E402
from test_oracle_kat import random_nodes, random_pods  # noqa: E402
from yoda_scheduler_b200 import api  # noqa: E402
from yoda_scheduler_b200.labels import pods_to_requests  # noqa: E402
from yoda_scheduler_b200.snapshot import Snapshot  # noqa: E402

first = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
count = int(sys.argv[2]) if len(sys.argv) > 2 else 300
paths = {0: 0, 1: 0}
for seed in range(first, first + count):
    for adversarial in (False, True):
        rng = random.Random(77 * seed + adversarial)
        nodes = random_nodes(rng, rng.choice([1, 3, 31, 32, 33, 127, 129, 300, 700, 1500]), adversarial)
        labels = random_pods(rng, rng.choice([1, 2, 9, 17, 129, 400, 1200]), adversarial)
        snap = Snapshot.from_nodes(nodes)
        for flags in (api.EVAL_FULL, api.EVAL_FULL | api.EVAL_FORCE_DIRECT):
            res, mats, t = gpu_cycle(snap, pods_to_requests(labels), flags)
            paths[t.path] += 1
            assert_matches_oracle(snap, labels, res, mats, what=f"seed={seed} adv={adversarial} flags={flags}")
# heterogeneous cards, many requirement groups and classes: always the table path
clocks = [877, 1215, 1410, 5705, 7000]
for seed in range(first, first + max(1, count // 5)):
    rng = random.Random(5000 + seed)
    nodes = []
    for i in range(rng.choice([200, 700, 1300])):
        nc = rng.choice([1, 2, 4, 8, 8])
        cards = []
        for g in range(nc):
            total = rng.choice([8000, 11000, 16000, 24000, 40536, 81920])
            cards.append({"health": "Healthy" if rng.random() > 0.08 else "Unhealthy", "free": rng.randrange(total + 1), "total": total,
                          "clock": rng.choice(clocks), "core": rng.choice([1, 2560, 3584, 4352, 5120, 6912]),
                          "bandwidth": rng.choice([1, 320, 352, 448, 616, 900, 1555]), "power": rng.choice([1, 70, 250, 260, 300, 400])})
        running = [{"scv/memory": str(rng.randrange(100, 30000))} for _ in range(rng.randrange(0, 3))]
        nodes.append({"name": f"n{i}", "has_scv": rng.random() > 0.02, "card_number": rng.choice([nc, nc, nc, 8, 0]), "cards": cards,
                      "free_sum": sum(c["free"] for c in cards), "total_sum": sum(c["total"] for c in cards), "pods": running})
    labels = []
    for _ in range(rng.choice([40, 600, 1500])):
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
    assert t.path == 1, t.path
    paths[1] += 1
    assert_matches_oracle(snap, labels, res, mats, what=f"heterogeneous seed={seed}")
print(f"soak ok: seeds {first}..{first + count - 1}, evaluations on the table path {paths[1]}, on the exact path {paths[0]}")
