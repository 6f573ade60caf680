"""Host-side structure-of-arrays image of the SCV cache (one row per node).

This is the host mirror from which the device snapshot can always be rebuilt; it also
converts to the array-of-structs forms the oracle and the pure-Python cross-check want.
Field meaning follows scv.Scv / scv.Card as used by the reference.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Mapping, Sequence

import numpy as np

G = 8  # YODA_MAX_CARDS

# binary layouts of include/yoda_b200.h
NODE_REC = np.dtype([
    ("card_number", "<u8"), ("free_memory_sum", "<u8"), ("total_memory_sum", "<u8"),
    ("n_cards", "<u4"), ("healthy_mask", "<u4"),
    ("free_memory", "<u8", (G,)), ("total_memory", "<u8", (G,)), ("clock", "<u8", (G,)),
    ("core", "<u8", (G,)), ("bandwidth", "<u8", (G,)), ("power", "<u8", (G,)),
])
POD_REQ = np.dtype([("flags", "<u4"), ("reserved", "<u4"), ("number", "<u8"), ("memory", "<u8"), ("clock", "<u8")])
POD_RES = np.dtype([
    ("max_bandwidth", "<u8"), ("max_clock", "<u8"), ("max_core", "<u8"), ("max_free_memory", "<u8"),
    ("max_power", "<u8"), ("max_total_memory", "<u8"), ("lowest", "<i8"), ("highest", "<i8"),
    ("feasible", "<u4"), ("chosen", "<i4"),
])
assert NODE_REC.itemsize == 416 and POD_REQ.itemsize == 32 and POD_RES.itemsize == 72

CARD_FIELDS = ("free_memory", "total_memory", "clock", "core", "bandwidth", "power")


@dataclass
class Snapshot:
    n_nodes: int
    has_scv: np.ndarray          # u8 [N]
    card_number: np.ndarray      # u64 [N]   Status.CardNumber
    n_cards: np.ndarray          # u8 [N]    len(Status.CardList)
    healthy_mask: np.ndarray     # u8 [N]    bit g: CardList[g].Health == "Healthy"
    free_memory: np.ndarray      # u64 [N, 8]
    total_memory: np.ndarray
    clock: np.ndarray
    core: np.ndarray
    bandwidth: np.ndarray
    power: np.ndarray
    free_memory_sum: np.ndarray  # u64 [N]
    total_memory_sum: np.ndarray
    pod_off: np.ndarray          # u32 [N+1] CSR over the running pods that carry scv/memory
    pod_mem: np.ndarray          # u64 [pod_off[-1]] their StrToUint64(scv/memory)
    names: List[str] = field(default_factory=list)

    @classmethod
    def empty(cls, n: int) -> "Snapshot":
        z64 = lambda *s: np.zeros(s, np.uint64)
        return cls(n, np.zeros(n, np.uint8), z64(n), np.zeros(n, np.uint8), np.zeros(n, np.uint8), z64(n, G), z64(n, G), z64(n, G),
                   z64(n, G), z64(n, G), z64(n, G), z64(n), np.ones(n, np.uint64), np.zeros(n + 1, np.uint32), z64(0),
                   [f"node-{i}" for i in range(n)])

    @classmethod
    def from_nodes(cls, nodes: Sequence[Mapping]) -> "Snapshot":
        """nodes: dicts as used by oracle/pyref.py (cards with string Health, pods with label maps)."""
        from .labels import LABEL_MEMORY, str_to_uint64
        s = cls.empty(len(nodes))
        off, mem = [0], []
        for i, nd in enumerate(nodes):
            s.has_scv[i] = 1 if nd.get("has_scv", True) else 0
            s.card_number[i] = nd["card_number"]
            cards = nd["cards"]
            assert len(cards) <= G
            s.n_cards[i] = len(cards)
            hm = 0
            for g, c in enumerate(cards):
                hm |= (1 << g) if c["health"] == "Healthy" else 0  # filter.go:53,57
                s.free_memory[i, g] = c["free"]
                s.total_memory[i, g] = c["total"]
                s.clock[i, g] = c["clock"]
                s.core[i, g] = c["core"]
                s.bandwidth[i, g] = c["bandwidth"]
                s.power[i, g] = c["power"]
            s.healthy_mask[i] = hm
            s.free_memory_sum[i] = nd["free_sum"]
            s.total_memory_sum[i] = nd["total_sum"]
            for lab in nd.get("pods", []):
                if LABEL_MEMORY in lab:  # algorithm.go:77
                    mem.append(str_to_uint64(lab[LABEL_MEMORY]))
            off.append(len(mem))
            if "name" in nd:
                s.names[i] = nd["name"]
        s.pod_off = np.array(off, np.uint32)
        s.pod_mem = np.array(mem, np.uint64)
        return s

    def alloc_sum(self) -> np.ndarray:
        """allocateMemorySum per node (algorithm.go:75-80), wrapping uint64."""
        cs = np.concatenate([np.zeros(1, np.uint64), np.cumsum(self.pod_mem, dtype=np.uint64)])
        return (cs[self.pod_off[1:]] - cs[self.pod_off[:-1]]).astype(np.uint64)

    def records(self, idx=None) -> np.ndarray:
        """yoda_node_record array for the given node indices (default: all)."""
        sel = np.arange(self.n_nodes) if idx is None else np.asarray(idx)
        r = np.zeros(len(sel), NODE_REC)
        r["card_number"] = self.card_number[sel]
        r["free_memory_sum"] = self.free_memory_sum[sel]
        r["total_memory_sum"] = self.total_memory_sum[sel]
        r["n_cards"] = self.n_cards[sel]
        r["healthy_mask"] = self.healthy_mask[sel]
        for f in CARD_FIELDS:
            r[f] = getattr(self, f)[sel]
        return r

    def to_nodes(self) -> list:
        """Array-of-structs view for oracle/pyref.py (small snapshots only)."""
        out = []
        for i in range(self.n_nodes):
            cards = [{"health": "Healthy" if (int(self.healthy_mask[i]) >> g) & 1 else "Unhealthy",
                      "free": int(self.free_memory[i, g]), "total": int(self.total_memory[i, g]), "clock": int(self.clock[i, g]),
                      "core": int(self.core[i, g]), "bandwidth": int(self.bandwidth[i, g]), "power": int(self.power[i, g])}
                     for g in range(int(self.n_cards[i]))]
            pods = [{"scv/memory": str(int(m) - (1 << 64) if int(m) >> 63 else int(m))} for m in self.pod_mem[self.pod_off[i]:self.pod_off[i + 1]]]
            out.append({"name": self.names[i] if self.names else f"node-{i}", "has_scv": bool(self.has_scv[i]),
                        "card_number": int(self.card_number[i]), "cards": cards, "free_sum": int(self.free_memory_sum[i]),
                        "total_sum": int(self.total_memory_sum[i]), "pods": pods})
        return out

    def slice(self, lo: int, hi: int) -> "Snapshot":
        """Nodes [lo, hi) as their own snapshot (node sharding)."""
        po = self.pod_off[lo:hi + 1]
        return Snapshot(hi - lo, self.has_scv[lo:hi], self.card_number[lo:hi], self.n_cards[lo:hi], self.healthy_mask[lo:hi],
                        self.free_memory[lo:hi], self.total_memory[lo:hi], self.clock[lo:hi], self.core[lo:hi], self.bandwidth[lo:hi],
                        self.power[lo:hi], self.free_memory_sum[lo:hi], self.total_memory_sum[lo:hi], (po - po[0]).astype(np.uint32),
                        self.pod_mem[po[0]:po[-1]], self.names[lo:hi] if self.names else [])
