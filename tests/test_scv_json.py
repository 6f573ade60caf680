"""SCV custom resource JSON -> yoda_node_record (host only, no GPU): the ingest step of SURVEY.md 8(f)4."""
import ctypes as C
import json

import numpy as np
import pytest

from yoda_scheduler_b200 import _lib
from yoda_scheduler_b200.snapshot import NODE_REC


def _parse(doc):
    L = _lib.load()
    text = doc if isinstance(doc, bytes) else json.dumps(doc).encode()
    rec = np.zeros(1, NODE_REC)
    name = C.create_string_buffer(64)
    rc = L.yoda_record_from_scv_json(text, len(text), rec.ctypes.data, name, 64)
    return rc, rec[0], name.value.decode()


CARD = {"id": 0, "health": "Healthy", "model": "A100", "power": 400, "totalMemory": 40536, "clock": 1215, "freeMemory": 40000, "core": 6912, "bandwidth": 1555}


def test_whole_object_and_status_alone_and_key_case():
    status = {"cardList": [CARD, dict(CARD, health="Unhealthy", freeMemory=2000), dict(CARD, freeMemory=18446744073709551615)],
              "cardNumber": 4, "freeMemorySum": 83436, "totalMemorySum": 162144, "updateTime": "2020-09-01T02:28:03Z"}
    for doc in ({"apiVersion": "core.run-linux.com/v1", "kind": "Scv", "metadata": {"name": "node-7", "labels": {"a": "b"}}, "spec": {}, "status": status},
                status,
                {"Status": {"CardList": status["cardList"], "CardNumber": 4, "FreeMemorySum": 83436, "TotalMemorySum": 162144}, "metadata": {"name": "node-7"}}):
        rc, r, name = _parse(doc)
        assert rc == 0
        assert name == ("node-7" if "metadata" in doc else "")
        assert (r["card_number"], r["free_memory_sum"], r["total_memory_sum"], r["n_cards"]) == (4, 83436, 162144, 3)
        assert r["healthy_mask"] == 0b101                      # exactly the string "Healthy" (filter.go:53,57)
        assert r["free_memory"][:3].tolist() == [40000, 2000, 2**64 - 1]
        assert r["total_memory"][:3].tolist() == [40536] * 3 and r["clock"][0] == 1215 and r["core"][1] == 6912
        assert r["bandwidth"][2] == 1555 and r["power"][0] == 400 and r["free_memory"][3:].sum() == 0


def test_edge_cases():
    rc, r, _ = _parse({"status": {"cardList": None, "cardNumber": 0, "freeMemorySum": 0, "totalMemorySum": 1}})
    assert rc == 0 and r["n_cards"] == 0 and r["total_memory_sum"] == 1
    rc, r, _ = _parse({"status": {"cardList": [], "totalMemorySum": 5}})
    assert rc == 0 and r["n_cards"] == 0
    rc, r, _ = _parse({"status": {"cardList": [dict(CARD, health="healthy")] * 8, "cardNumber": 8, "totalMemorySum": 9}})
    assert rc == 0 and r["n_cards"] == 8 and r["healthy_mask"] == 0          # lower-case "healthy" is not "Healthy"
    rc, r, _ = _parse({"status": {"cardList": [CARD] * 9, "cardNumber": 9, "totalMemorySum": 9}})
    assert rc == -3 and r["n_cards"] == 9                                      # YODA_E_RANGE
    for bad in (b"", b"[]", b'{"status": {"cardNumber": -1}}', b'{"status": {"cardNumber": 1.5}}', b'{"status": {"cardList": [{"health": 3}]}}',
                b'{"status": {"cardNumber": 18446744073709551616}}', b'{"status": {"cardList": [', b'{"metadata": {"name": "x"}}', b'{"status": {"other": 1}}'):
        assert _parse(bad)[0] == -1, bad                                       # YODA_E_ARG


@pytest.mark.gpu
def test_json_records_feed_the_snapshot():
    """Records parsed from JSON behave exactly like records built from the same numbers."""
    from oracle import oracle as ora
    from parity import assert_matches_oracle
    from yoda_scheduler_b200 import api, synth
    from yoda_scheduler_b200.labels import pods_to_requests
    snap = synth.make_snapshot(300)
    nodes = snap.to_nodes()
    labels = synth.make_pod_labels(120, 21)
    with api.Context(300) as ctx:
        for i, nd in enumerate(nodes):
            if not nd["has_scv"]:
                ctx.remove(i)
                continue
            doc = {"metadata": {"name": nd["name"]}, "status": {"cardNumber": nd["card_number"], "freeMemorySum": nd["free_sum"], "totalMemorySum": nd["total_sum"],
                   "cardList": [{"health": c["health"], "freeMemory": c["free"], "totalMemory": c["total"], "clock": c["clock"], "core": c["core"],
                                 "bandwidth": c["bandwidth"], "power": c["power"]} for c in nd["cards"]]}}
            rc, rec, name = _parse(doc)
            assert rc == 0 and name == nd["name"]
            ctx.upsert(i, rec)
        ctx.set_alloc_many(np.arange(300), snap.alloc_sum())
        ctx.commit()
        res = ctx.eval(pods_to_requests(labels), api.EVAL_FULL)
        assert_matches_oracle(snap, labels, res, ctx.fetch_all(len(labels)))
