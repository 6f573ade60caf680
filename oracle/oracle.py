"""ctypes binding of the CPU oracle (oracle/yoda_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libyoda_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "yoda_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.ora_cluster_new.restype = C.c_void_p
        L.ora_cluster_new.argtypes = [C.c_uint32]
        L.ora_cluster_free.argtypes = [C.c_void_p]
        L.ora_str_to_uint64.restype = C.c_uint64
        L.ora_str_to_uint64.argtypes = [C.c_char_p]
        L.ora_uint64_to_int64.restype = C.c_int64
        L.ora_uint64_to_int64.argtypes = [C.c_uint64]
        L.ora_normalize_score.argtypes = [C.c_void_p, C.c_uint32]
        L.ora_cluster_set_node.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_uint64, C.c_uint32] + [C.c_void_p] * 7 + [C.c_uint64, C.c_uint64]
        L.ora_cluster_set_node_pods.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ora_cluster_fill.argtypes = [C.c_void_p] * 15
        L.ora_eval_batch.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 8
        L.ora_filter.argtypes = [C.c_void_p, C.c_uint32, C.c_char_p, C.c_char_p, C.c_char_p]
        L.ora_node_static.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.ora_score_with_max.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ora_eval_sequential.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 9
        L.ora_pod_priority.restype = C.c_int64
        L.ora_pod_priority.argtypes = [C.c_char_p]
        L.ora_queue_order.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p]
        L.ora_num_feasible_nodes_to_find.restype = C.c_uint32
        L.ora_num_feasible_nodes_to_find.argtypes = [C.c_uint32, C.c_int32]
        L.ora_framework_tail.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int32, C.c_uint64, C.c_uint32,
                                         C.c_int] + [C.c_void_p] * 9
        _lib = L
    return _lib


def str_to_uint64(s: str) -> int:
    return int(lib().ora_str_to_uint64(s.encode("utf-8", "surrogateescape")))


def uint64_to_int64(u: int) -> int:
    return int(lib().ora_uint64_to_int64(C.c_uint64(u & (2**64 - 1))))


def normalize_score(scores: Sequence[int]) -> list:
    a = np.array([s for s in scores], dtype=np.int64)
    rc = lib().ora_normalize_score(a.ctypes.data, len(a))
    if rc:
        raise IndexError("NormalizeScore on an empty list panics in the reference (scheduler.go:135)")
    return a.tolist()


def num_feasible_nodes_to_find(n_nodes: int, percentage: int = 0) -> int:
    return int(lib().ora_num_feasible_nodes_to_find(n_nodes, percentage))


def pod_priority(label: Optional[str]) -> int:
    return int(lib().ora_pod_priority(None if label is None else label.encode()))


def queue_order(priority_labels: Sequence[Optional[str]]) -> np.ndarray:
    """Stable batch order under sort.Less (sort.go:8-18)."""
    arr = _cstr_array(list(priority_labels))
    order = np.zeros(len(priority_labels), np.uint32)
    lib().ora_queue_order(len(priority_labels), C.cast(arr, C.c_void_p), order.ctypes.data)
    return order


def _cstr_array(vals: Sequence[Optional[str]]):
    arr = (C.c_char_p * max(len(vals), 1))()
    for i, v in enumerate(vals):
        arr[i] = None if v is None else v.encode("utf-8", "surrogateescape")
    return arr


def _p(a):
    return None if a is None else a.ctypes.data


class OracleCluster:
    """AoS copy of a snapshot inside the C oracle."""

    def __init__(self, n_nodes: int):
        self.n = int(n_nodes)
        self._h = C.c_void_p(lib().ora_cluster_new(self.n))

    def __del__(self):
        try:
            if self._h:
                lib().ora_cluster_free(self._h)
                self._h = None
        except Exception:
            pass

    def set_node(self, idx, has_scv, card_number, health, free, total, clock, core, bandwidth, power, free_sum, total_sum):
        n = len(health)
        u = lambda xs: np.array([x & (2**64 - 1) for x in xs], dtype=np.uint64)
        arrs = [u(free), u(total), u(clock), u(core), u(bandwidth), u(power)]
        hs = _cstr_array(list(health))
        rc = lib().ora_cluster_set_node(self._h, idx, int(bool(has_scv)), C.c_uint64(card_number & (2**64 - 1)), n, C.cast(hs, C.c_void_p),
                                        *[a.ctypes.data for a in arrs], C.c_uint64(free_sum & (2**64 - 1)), C.c_uint64(total_sum & (2**64 - 1)))
        assert rc == 0

    def set_node_pods(self, idx, pods_labels: Sequence[dict]):
        nkv = np.array([len(d) for d in pods_labels], dtype=np.uint32)
        keys, vals = [], []
        for d in pods_labels:
            for k, v in d.items():
                keys.append(k)
                vals.append(v)
        ka, va = _cstr_array(keys), _cstr_array(vals)
        rc = lib().ora_cluster_set_node_pods(self._h, idx, len(pods_labels), nkv.ctypes.data if len(nkv) else None, C.cast(ka, C.c_void_p), C.cast(va, C.c_void_p))
        assert rc == 0

    @classmethod
    def from_snapshot(cls, snap) -> "OracleCluster":
        """snap: yoda_scheduler_b200.snapshot.Snapshot (numpy SoA)."""
        self = cls(snap.n_nodes)
        c64 = lambda a: np.ascontiguousarray(a, dtype=np.uint64)
        c8 = lambda a: np.ascontiguousarray(a, dtype=np.uint8)
        keep = [c8(snap.has_scv), c64(snap.card_number), c8(snap.n_cards), c8(snap.healthy_mask), c64(snap.free_memory), c64(snap.total_memory),
                c64(snap.clock), c64(snap.core), c64(snap.bandwidth), c64(snap.power), c64(snap.free_memory_sum), c64(snap.total_memory_sum),
                np.ascontiguousarray(snap.pod_off, dtype=np.uint32), c64(snap.pod_mem) if len(snap.pod_mem) else np.zeros(1, np.uint64)]
        rc = lib().ora_cluster_fill(self._h, *[a.ctypes.data for a in keep])
        assert rc == 0
        return self

    def filter(self, node: int, labels: dict) -> bool:
        enc = lambda k: None if k not in labels else labels[k].encode()
        return bool(lib().ora_filter(self._h, node, enc("scv/number"), enc("scv/memory"), enc("scv/clock")))

    def node_static(self, node: int):
        a, b = C.c_uint64(0), C.c_uint64(0)
        rc = lib().ora_node_static(self._h, node, C.byref(a), C.byref(b))
        if rc:
            raise ZeroDivisionError("TotalMemorySum == 0 panics in the reference (algorithm.go:71,86)")
        return a.value, b.value

    def score_with_max(self, labels: dict, max6):
        """(feasible, raw) of every node for one pod under given collection.MaxValue maxima."""
        enc = lambda k: None if k not in labels else labels[k].encode()
        mx = np.ascontiguousarray(max6, dtype=np.uint64)
        feas = np.zeros(self.n, np.uint8)
        raw = np.zeros(self.n, np.int64)
        rc = lib().ora_score_with_max(self._h, enc("scv/number"), enc("scv/memory"), enc("scv/clock"), mx.ctypes.data, feas.ctypes.data, raw.ctypes.data)
        assert rc == 0, rc
        return feas, raw

    def framework_tail(self, pods: Sequence[dict], weight: int = 300, percentage: int = 0, seed: int = 0, start_index: int = 0, chain: bool = False,
                       other=None) -> dict:
        """What kube-scheduler v1.20 does around the plugin per pod: sample feasible nodes round-robin, score + normalize the sample,
        weight, add the other plugins' scores, seeded reservoir tie-break (ora_framework_tail)."""
        P = len(pods)
        num = _cstr_array([d.get("scv/number") for d in pods])
        mem = _cstr_array([d.get("scv/memory") for d in pods])
        clk = _cstr_array([d.get("scv/clock") for d in pods])
        o = None if other is None else np.ascontiguousarray(other, dtype=np.int64)
        out = {"chosen": np.zeros(P, np.int32), "sampled": np.zeros(P, np.uint32), "processed": np.zeros(P, np.uint32), "start": np.zeros(P, np.uint32),
               "lowest": np.zeros(P, np.int64), "highest": np.zeros(P, np.int64), "top": np.zeros(P, np.int64), "ties": np.zeros(P, np.uint32)}
        rc = lib().ora_framework_tail(self._h, P, C.cast(num, C.c_void_p), C.cast(mem, C.c_void_p), C.cast(clk, C.c_void_p), weight, percentage,
                                      C.c_uint64(seed & (2**64 - 1)), start_index, int(chain), _p(o), *[_p(out[k]) for k in
                                      ("chosen", "sampled", "processed", "start", "lowest", "highest", "top", "ties")])
        assert rc == 0, rc
        return out

    def eval_sequential(self, pods: Sequence[dict], order=None) -> dict:
        """Single-pod cycles in `order`, each pod assumed onto its chosen node before the next one."""
        P = len(pods)
        num = _cstr_array([d.get("scv/number") for d in pods])
        mem = _cstr_array([d.get("scv/memory") for d in pods])
        clk = _cstr_array([d.get("scv/clock") for d in pods])
        o = None if order is None else np.ascontiguousarray(order, dtype=np.uint32)
        out = {"max6": np.zeros((P, 6), np.uint64), "lowest": np.zeros(P, np.int64), "highest": np.zeros(P, np.int64),
               "count": np.zeros(P, np.uint32), "chosen": np.zeros(P, np.int32)}
        rc = lib().ora_eval_sequential(self._h, P, C.cast(num, C.c_void_p), C.cast(mem, C.c_void_p), C.cast(clk, C.c_void_p), _p(o),
                                       _p(out["max6"]), _p(out["lowest"]), _p(out["highest"]), _p(out["count"]), _p(out["chosen"]))
        assert rc == 0, rc
        return out

    def eval(self, pods: Sequence[dict], faithful: bool = False, threads: int = 1, full: bool = True) -> dict:
        """pods: list of label dicts (keys scv/number, scv/memory, scv/clock; others ignored)."""
        P, N = len(pods), self.n
        num = _cstr_array([d.get("scv/number") for d in pods])
        mem = _cstr_array([d.get("scv/memory") for d in pods])
        clk = _cstr_array([d.get("scv/clock") for d in pods])
        out = {
            "feasible": np.zeros((P, N), np.uint8) if full else None,
            "raw": np.zeros((P, N), np.int64) if full else None,
            "norm": np.zeros((P, N), np.int64) if full else None,
            "max6": np.zeros((P, 6), np.uint64),
            "lowest": np.zeros(P, np.int64),
            "highest": np.zeros(P, np.int64),
            "count": np.zeros(P, np.uint32),
            "chosen": np.zeros(P, np.int32),
        }
        rc = lib().ora_eval_batch(self._h, P, C.cast(num, C.c_void_p), C.cast(mem, C.c_void_p), C.cast(clk, C.c_void_p), int(faithful), int(threads),
                                  _p(out["feasible"]), _p(out["raw"]), _p(out["norm"]), _p(out["max6"]), _p(out["lowest"]), _p(out["highest"]),
                                  _p(out["count"]), _p(out["chosen"]))
        if rc == -2:
            raise ZeroDivisionError("TotalMemorySum == 0 panics in the reference (algorithm.go:71,86)")
        assert rc == 0, rc
        return out
