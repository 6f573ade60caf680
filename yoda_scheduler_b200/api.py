"""Context: Python handle on one yoda_ctx (one GPU, one snapshot shard)."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib
from ._lib import Config, Timing, Xchg, YodaError, check  # noqa: F401
from .snapshot import NODE_REC, POD_REQ, POD_RES, Snapshot

EVAL_DECIDE, EVAL_FULL, EVAL_FORCE_DIRECT, EVAL_RAW32 = 0, 1, 2, 4
ROW_FEASIBLE_BITS, ROW_RAW, ROW_NORM = 0, 1, 2
TAIL_RES = np.dtype([("chosen", np.int32), ("sampled", np.uint32), ("processed", np.uint32), ("start", np.uint32), ("lowest", np.int64),
                     ("highest", np.int64), ("top", np.int64), ("ties", np.uint32), ("reserved", np.uint32)])


class Context:
    def __init__(self, max_nodes: int, device: int = 0, node_base: int = 0, shard_rank: int = 0, shard_count: int = 1):
        self.L = _lib.load()
        cfg = Config(device, max_nodes, node_base, shard_rank, shard_count, 0)
        h = C.c_void_p()
        check(self.L.yoda_create(C.byref(cfg), C.byref(h)))
        self.h = h
        self.max_nodes = max_nodes
        self.node_base = node_base

    def close(self):
        if getattr(self, "h", None):
            self.L.yoda_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- snapshot maintenance ------------------------------------------------------------
    def upsert(self, idx: int, rec: np.ndarray):
        rec = np.ascontiguousarray(rec, dtype=NODE_REC).reshape(1)
        check(self.L.yoda_node_upsert(self.h, idx, rec.ctypes.data))

    def upsert_many(self, idx, recs: np.ndarray):
        idx = np.ascontiguousarray(idx, dtype=np.uint32)
        recs = np.ascontiguousarray(recs, dtype=NODE_REC)
        assert len(idx) == len(recs)
        check(self.L.yoda_node_upsert_many(self.h, len(idx), idx.ctypes.data, recs.ctypes.data))

    def set_alloc(self, idx: int, alloc: int):
        check(self.L.yoda_node_set_alloc(self.h, idx, C.c_uint64(alloc & (2**64 - 1))))

    def set_alloc_many(self, idx, alloc):
        idx = np.ascontiguousarray(idx, dtype=np.uint32)
        alloc = np.ascontiguousarray(alloc, dtype=np.uint64)
        check(self.L.yoda_node_set_alloc_many(self.h, len(idx), idx.ctypes.data, alloc.ctypes.data))

    def remove(self, idx: int):
        check(self.L.yoda_node_remove(self.h, idx))

    def commit(self) -> int:
        e = C.c_uint64()
        check(self.L.yoda_commit(self.h, C.byref(e)))
        return e.value

    def node_count(self) -> int:
        n = C.c_uint32()
        check(self.L.yoda_node_count(self.h, C.byref(n)))
        return n.value

    def load_snapshot(self, snap: Snapshot, commit: bool = True):
        """Full (re)build of the device snapshot from the host mirror."""
        idx = np.arange(snap.n_nodes, dtype=np.uint32)
        have = snap.has_scv.astype(bool)
        if have.any():
            self.upsert_many(idx[have], snap.records(idx[have]))
        for i in idx[~have]:
            self.remove(int(i))
        self.set_alloc_many(idx, snap.alloc_sum())
        if commit:
            return self.commit()

    # ---- evaluation ------------------------------------------------------------------------
    def eval(self, reqs: np.ndarray, flags: int = EVAL_FULL) -> np.ndarray:
        reqs = np.ascontiguousarray(reqs, dtype=POD_REQ)
        out = np.zeros(len(reqs), POD_RES)
        check(self.L.yoda_eval_batch(self.h, reqs.ctypes.data, len(reqs), flags, out.ctypes.data))
        self._n = None
        return out

    def eval_device(self, d_pods_ptr: int, n_pods: int, flags: int = EVAL_FULL, d_out_ptr: Optional[int] = None):
        check(self.L.yoda_eval_batch_device(self.h, C.c_void_p(d_pods_ptr), n_pods, flags, C.c_void_p(d_out_ptr) if d_out_ptr else None))

    def sync(self):
        check(self.L.yoda_sync(self.h))

    def timing(self) -> Timing:
        t = Timing()
        check(self.L.yoda_last_timing(self.h, C.byref(t)))
        return t

    def fetch_row(self, pod: int, kind: int, n_nodes: Optional[int] = None) -> np.ndarray:
        n = self.node_count() if n_nodes is None else n_nodes
        if kind == ROW_FEASIBLE_BITS:
            words = np.zeros((n + 31) // 32, np.uint32)
            check(self.L.yoda_fetch_row(self.h, pod, kind, words.ctypes.data, words.nbytes))
            return np.unpackbits(words.view(np.uint8), bitorder="little")[:n].astype(np.uint8)
        out = np.zeros(n, np.int64 if kind == ROW_RAW else np.uint8)
        check(self.L.yoda_fetch_row(self.h, pod, kind, out.ctypes.data, out.nbytes))
        return out

    def result_raw_bytes(self) -> int:
        """4 when the last FULL result holds uint32 scores (EVAL_RAW32 honoured), else 8."""
        b = C.c_uint32()
        check(self.L.yoda_result_raw_bytes(self.h, C.byref(b)))
        return b.value

    def fetch_all(self, n_pods: int) -> dict:
        n = self.node_count()
        return {
            "feasible": np.stack([self.fetch_row(p, ROW_FEASIBLE_BITS, n) for p in range(n_pods)]) if n_pods else np.zeros((0, n), np.uint8),
            "raw": np.stack([self.fetch_row(p, ROW_RAW, n) for p in range(n_pods)]) if n_pods else np.zeros((0, n), np.int64),
            "norm": np.stack([self.fetch_row(p, ROW_NORM, n) for p in range(n_pods)]) if n_pods else np.zeros((0, n), np.uint8),
        }

    def pod_order(self, n_pods: int):
        """(orig, group, cls, score_chunk, out_chunk) of the last table-path evaluation (yoda_debug_pod_order)."""
        orig, grp, cls = (np.zeros(n_pods, np.uint32) for _ in range(3))
        sc, oc = C.c_uint32(), C.c_uint32()
        check(self.L.yoda_debug_pod_order(self.h, orig.ctypes.data, grp.ctypes.data, cls.ctypes.data, n_pods, C.byref(sc), C.byref(oc)))
        return orig, grp, cls, sc.value, oc.value

    def framework_tail(self, n_pods: int, weight: int = 300, percentage: int = 0, seed: int = 0, start_index: int = 0, chain: bool = False, other=None):
        """Sampling, weight and seeded selectHost over the last FULL batch (yoda_framework_tail).  Returns (results, next_start)."""
        out = np.zeros(n_pods, TAIL_RES)
        o = None if other is None else np.ascontiguousarray(other, dtype=np.int64)
        cfg = _lib.TailConfig(weight, percentage, seed & (2**64 - 1), start_index, 1 if chain else 0, None if o is None else o.ctypes.data)
        nxt = C.c_uint32()
        check(self.L.yoda_framework_tail(self.h, C.byref(cfg), out.ctypes.data, C.byref(nxt)))
        return out, nxt.value

    def normalize_scores(self, scores) -> np.ndarray:
        a = np.ascontiguousarray(scores, dtype=np.int64).copy()
        check(self.L.yoda_normalize_scores(self.h, a.ctypes.data, len(a)))
        return a

    def eval_sequential(self, reqs: np.ndarray, order=None) -> np.ndarray:
        """One pod after the other, each choice committed to the chosen node's allocateMemorySum first."""
        reqs = np.ascontiguousarray(reqs, dtype=POD_REQ)
        out = np.zeros(len(reqs), POD_RES)
        o = None if order is None else np.ascontiguousarray(order, dtype=np.uint32)
        check(self.L.yoda_eval_sequential(self.h, reqs.ctypes.data, len(reqs), None if o is None else o.ctypes.data, out.ctypes.data))
        return out

    def queue_order(self, priorities) -> np.ndarray:
        """Batch consumption order under sort.Less (scv/priority, higher first, stable)."""
        pri = np.ascontiguousarray(priorities, dtype=np.int64)
        order = np.zeros(len(pri), np.uint32)
        check(self.L.yoda_queue_order(self.h, pri.ctypes.data, len(pri), order.ctypes.data))
        return order

    # ---- phase-split evaluation (host-driven exchange between node shards) ---------------
    def phase_begin(self, reqs: np.ndarray, flags: int = EVAL_FULL):
        reqs = np.ascontiguousarray(reqs, dtype=POD_REQ)
        self._ph_n = len(reqs)
        check(self.L.yoda_phase_begin(self.h, reqs.ctypes.data, len(reqs), flags))

    def phase_run(self, phase: int):
        check(self.L.yoda_phase_run(self.h, phase))

    def phase_exchange(self, phase: int):
        bufs = (Xchg * 4)()
        n = C.c_int(4)
        check(self.L.yoda_phase_exchange(self.h, phase, bufs, C.byref(n)))
        return [(bufs[i].d_ptr, bufs[i].count, bufs[i].op) for i in range(n.value)]

    def phase_finish(self) -> np.ndarray:
        out = np.zeros(self._ph_n, POD_RES)
        check(self.L.yoda_phase_finish(self.h, out.ctypes.data))
        return out


def parse_label(s: str) -> int:
    b = s.encode("utf-8", "surrogateescape")
    return int(_lib.load().yoda_parse_label(b, len(b)))
