import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


@pytest.fixture(scope="session")
def kat():
    with open(os.path.join(GOLDEN, "kat_cluster.json")) as f:
        return json.load(f)


def kat_nodes(kat):
    """KAT cluster as the node dicts of oracle/pyref.py."""
    keys = ("health", "free", "total", "clock", "core", "bandwidth", "power")
    out = []
    for nd in kat["nodes"]:
        out.append({"name": nd["name"], "has_scv": nd["has_scv"], "card_number": nd["card_number"], "free_sum": nd["free_sum"],
                    "total_sum": nd["total_sum"], "cards": [dict(zip(keys, c)) for c in nd["cards"]], "pods": nd["pods"]})
    return out


def oracle_cluster_from_nodes(nodes):
    """Build the C oracle's AoS cluster from node dicts, keeping the literal strings."""
    from oracle.oracle import OracleCluster
    cl = OracleCluster(len(nodes))
    for i, nd in enumerate(nodes):
        cs = nd["cards"]
        cl.set_node(i, nd.get("has_scv", True), nd["card_number"], [c["health"] for c in cs], [c["free"] for c in cs],
                    [c["total"] for c in cs], [c["clock"] for c in cs], [c["core"] for c in cs], [c["bandwidth"] for c in cs],
                    [c["power"] for c in cs], nd["free_sum"], nd["total_sum"])
        cl.set_node_pods(i, nd.get("pods", []))
    return cl


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
