"""Synthetic SCV snapshots and pending-pod batches.

Counter-based SplitMix64: every value is a pure function of (seed, stream, i, j), so the C++
harness, Python and any other host generate identical data.  Vectorised numpy uint64
arithmetic (wraps modulo 2^64 like the C++ twin).
"""
from __future__ import annotations

from typing import List

import numpy as np

from .snapshot import G, Snapshot

SEED = 0x59304441
_GOLD = np.uint64(0x9E3779B97F4A7C15)
_C1 = np.uint64(0xBF58476D1CE4E5B9)
_C2 = np.uint64(0x94D049BB133111EB)

# streams
S_SKU, S_FREE, S_HEALTH, S_CARDNUM, S_NOSCV, S_NPODS, S_PODMEM, S_NCARDS = 1, 2, 3, 4, 5, 6, 7, 8
S_P_NUM, S_P_NUMV, S_P_MEM, S_P_MEMV, S_P_CLK, S_P_CLKV = 16, 17, 18, 19, 20, 21

# fixed SKU table: (total MB, clock, cores, bandwidth, power) -- homogeneous within a node
SKU_TABLE = np.array([
    (11178, 5705, 3584, 352, 250),     # GTX 1080 Ti class (clock value of readme.md:62)
    (11019, 7000, 4352, 616, 260),     # RTX 2080 Ti class
    (16160, 877, 5120, 900, 300),      # V100 16G class
    (32510, 877, 5120, 900, 300),      # V100 32G class
    (40536, 1215, 6912, 1555, 400),    # A100 40G class
    (81920, 1593, 6912, 2039, 400),    # A100 80G class
    (81559, 2619, 16896, 3350, 700),   # H100 class
    (183359, 3996, 18944, 8000, 1000), # B200 class
], dtype=np.uint64)
BOGUS_CLOCK = 1234  # a clock no card has


def _sm64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = x + _GOLD
        z = (z ^ (z >> np.uint64(30))) * _C1
        z = (z ^ (z >> np.uint64(27))) * _C2
        return z ^ (z >> np.uint64(31))


def rnd(seed: int, stream: int, i, j=0) -> np.ndarray:
    """r(seed, stream, i, j) = sm64(sm64(sm64(seed + stream) + i) + j)"""
    with np.errstate(over="ignore"):
        h = _sm64(np.asarray(np.uint64((seed + stream) & (2**64 - 1))))
        h = _sm64(h + np.asarray(i, dtype=np.uint64))
        return _sm64(h + np.asarray(j, dtype=np.uint64))


def make_snapshot(n_nodes: int, seed: int = SEED, node_offset: int = 0) -> Snapshot:
    """Nodes [node_offset, node_offset + n_nodes) of the synthetic cluster."""
    N = n_nodes
    idx = np.arange(node_offset, node_offset + N, dtype=np.uint64)
    card = np.arange(G, dtype=np.uint64)[None, :]
    s = Snapshot.empty(N)
    s.names = [f"node-{int(i)}" for i in idx] if N <= 100000 else []
    sku = SKU_TABLE[(rnd(seed, S_SKU, idx) % np.uint64(8)).astype(np.int64)]          # [N,5]
    ncards = np.full(N, G, np.uint64)
    odd = rnd(seed, S_NCARDS, idx)
    ncards = np.where(odd % np.uint64(512) == 0, (odd >> np.uint64(16)) % np.uint64(9), ncards)
    live = card < ncards[:, None]
    total = np.where(live, sku[:, 0:1], np.uint64(0))
    r = rnd(seed, S_FREE, idx[:, None], card) >> np.uint64(32)
    free = np.where(live, (total * r) >> np.uint64(32), np.uint64(0))
    healthy = (rnd(seed, S_HEALTH, idx[:, None], card) % np.uint64(64) != 0) & live
    s.free_memory[:] = free
    s.total_memory[:] = total
    for k, name in ((1, "clock"), (2, "core"), (3, "bandwidth"), (4, "power")):
        getattr(s, name)[:] = np.where(live, sku[:, k:k + 1], np.uint64(0))
    s.healthy_mask[:] = (healthy.astype(np.uint64) << card).sum(axis=1).astype(np.uint8)
    s.n_cards[:] = ncards.astype(np.uint8)
    cn = rnd(seed, S_CARDNUM, idx)
    s.card_number[:] = np.where(cn % np.uint64(128) == 0, (cn >> np.uint64(8)) % np.uint64(17), np.uint64(G))
    s.has_scv[:] = (rnd(seed, S_NOSCV, idx) % np.uint64(256) != 0).astype(np.uint8)
    s.free_memory_sum[:] = free.sum(axis=1, dtype=np.uint64)
    s.total_memory_sum[:] = np.maximum(total.sum(axis=1, dtype=np.uint64), np.uint64(1))
    npods = (rnd(seed, S_NPODS, idx) % np.uint64(9)).astype(np.int64)
    s.pod_off = np.concatenate([[0], np.cumsum(npods)]).astype(np.uint32)
    node_of = np.repeat(np.arange(N), npods)
    slot = np.arange(len(node_of)) - s.pod_off[:-1].astype(np.int64)[node_of]
    k = rnd(seed, S_PODMEM, idx[node_of], slot.astype(np.uint64)) % np.uint64(128) + np.uint64(1)
    s.pod_mem = (k * np.uint64(256)).astype(np.uint64)
    return s


def make_pod_fields(n_pods: int, seed: int, round_values: bool = False):
    """flags/number/memory/clock arrays of a pending batch (requirement vectors)."""
    p = np.arange(n_pods, dtype=np.uint64)
    u = lambda st: rnd(seed, st, p)
    has_num = u(S_P_NUM) % np.uint64(100) >= 50
    v = u(S_P_NUMV)
    rare = (v >> np.uint64(20)) % np.uint64(100) == 0
    number = np.where(rare, np.where(v & np.uint64(1), np.uint64(9), np.uint64(0)),
                      np.array([1, 2, 4, 8], np.uint64)[((v >> np.uint64(8)) % np.uint64(4)).astype(np.int64)])
    has_mem = u(S_P_MEM) % np.uint64(100) >= 20
    memory = u(S_P_MEMV) % np.uint64(90000) + np.uint64(1)
    if round_values:
        memory = (memory // np.uint64(1000) + np.uint64(1)) * np.uint64(1000)
    has_clk = u(S_P_CLK) % np.uint64(100) >= 85
    v = u(S_P_CLKV)
    clock = np.where((v >> np.uint64(20)) % np.uint64(100) == 0, np.uint64(BOGUS_CLOCK),
                     SKU_TABLE[((v >> np.uint64(8)) % np.uint64(8)).astype(np.int64), 1])
    return has_num, number, has_mem, memory, has_clk, clock


def make_requests(n_pods: int, seed: int, round_values: bool = False) -> np.ndarray:
    from .snapshot import POD_REQ
    hn, num, hm, mem, hc, clk = make_pod_fields(n_pods, seed, round_values)
    r = np.zeros(n_pods, POD_REQ)
    r["flags"] = hn.astype(np.uint32) | (hm.astype(np.uint32) << 1) | (hc.astype(np.uint32) << 2)
    r["number"] = np.where(hn, num, 0)
    r["memory"] = np.where(hm, mem, 0)
    r["clock"] = np.where(hc, clk, 0)
    return r


def make_pod_labels(n_pods: int, seed: int, round_values: bool = False) -> List[dict]:
    """The same batch as label maps (what the Go plugin and the oracle see)."""
    hn, num, hm, mem, hc, clk = make_pod_fields(n_pods, seed, round_values)
    out = []
    for i in range(n_pods):
        d = {}
        if hn[i]:
            d["scv/number"] = str(int(num[i]))
        if hm[i]:
            d["scv/memory"] = str(int(mem[i]))
        if hc[i]:
            d["scv/clock"] = str(int(clk[i]))
        out.append(d)
    return out


# BASELINE.json configs: (nodes, pods, pod seed)
CONFIGS = {
    "c1": (100, 1, SEED + 1),
    "c2": (5000, 1000, SEED + 2),
    "c3": (50000, 100000, SEED + 3),
    "c4": (1000000, 10000, SEED + 4),
    "c5": (50000, 10000, SEED + 5),
}
