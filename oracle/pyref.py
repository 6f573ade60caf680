"""Second, independent restatement of the reference arithmetic in pure Python.

TEST INFRASTRUCTURE ONLY (small cases).  Arbitrary-precision ints masked to 64 bits stand in
for Go's uint/uint64; it exists to cross-check oracle/yoda_oracle.c, which in turn pins the
CUDA path.  PARITY UNPINNED by the reference itself (it has no tests); see yoda_oracle.c.

Citations are paths under /root/reference.
"""
from __future__ import annotations

M64 = (1 << 64) - 1
I63 = 1 << 63


def go_atoi(s: str):
    """[3P] strconv.Atoi: returns (value, ok)."""
    if not s:
        return 0, False
    body = s[1:] if s[0] in "+-" else s
    if not body or any(ch not in "0123456789" for ch in body):
        return 0, False
    v = int(body)
    if s[0] == "-":
        v = -v
    if v < -I63 or v > I63 - 1:
        return 0, False
    return v, True


def str_to_u64(s: str) -> int:
    """filter.go:60-74"""
    v, ok = go_atoi(s)
    return (v & M64) if ok else 0


def u64_to_i64(u: int) -> int:
    """filter.go:84-86"""
    v, ok = go_atoi(str(u & M64))
    return v if ok else 0


def fits_number(labels, node):
    """filter.go:11-16"""
    if "scv/number" in labels:
        n = str_to_u64(labels["scv/number"])
        return n <= node["card_number"], n
    return node["card_number"] > 0, 1


def fits_memory(number, labels, node):
    """filter.go:18-33,52-54"""
    if "scv/memory" in labels:
        m = str_to_u64(labels["scv/memory"])
        k = sum(1 for c in node["cards"] if c["health"] == "Healthy" and c["free"] >= m)
        return k >= number, m
    return True, 0


def fits_clock(number, labels, node):
    """filter.go:35-50,56-58"""
    if "scv/clock" in labels:
        ck = str_to_u64(labels["scv/clock"])
        k = sum(1 for c in node["cards"] if c["health"] == "Healthy" and c["clock"] == ck)
        return k >= number, ck
    return True, 0


def predicates(labels, node):
    ok, number = fits_number(labels, node)
    if not ok:
        return False, 0, 0
    fm, m = fits_memory(number, labels, node)
    fc, ck = fits_clock(number, labels, node)
    return fm and fc, m, ck


def yoda_filter(labels, node) -> bool:
    """scheduler.go:76-93"""
    if not node["has_scv"]:
        return False
    return predicates(labels, node)[0]


def collect_max(labels, nodes):
    """collection.go:30-78; order bw, clock, core, free, power, total (collection.go:14-21)"""
    mx = {"bandwidth": 1, "clock": 1, "core": 1, "free": 1, "power": 1, "total": 1}
    for node in nodes:
        if not node["has_scv"]:
            continue
        fit, m, ck = predicates(labels, node)
        if not fit:
            continue
        for c in node["cards"]:
            if c["free"] >= m and c["clock"] >= ck:
                for f in mx:
                    mx[f] = max(mx[f], c[f])
    return mx


def card_score(mx, c) -> int:
    """algorithm.go:57-68 (clock is divided by MaxBandwidth, :60)"""
    d = lambda x, y: ((x * 100) & M64) // y
    a = (d(c["bandwidth"], mx["bandwidth"]) + d(c["clock"], mx["bandwidth"]) + d(c["core"], mx["core"]) + d(c["power"], mx["power"])) & M64
    return (a + ((d(c["free"], mx["free"]) * 2) & M64) + d(c["total"], mx["total"])) & M64


def basic_score(mx, labels, node) -> int:
    """algorithm.go:41-55"""
    fit, m, ck = predicates(labels, node)
    if not fit:
        return 0
    s = 0
    for c in node["cards"]:
        if c["free"] >= m and c["clock"] >= ck:
            s = (s + card_score(mx, c)) & M64
    return s


def actual_score(node) -> int:
    """algorithm.go:70-72 (ZeroDivisionError == Go panic)"""
    return ((((node["free_sum"] * 100) & M64) // node["total_sum"]) * 2) & M64


def allocate_score(node) -> int:
    """algorithm.go:74-87"""
    a = 0
    for labels in node.get("pods", []):
        if "scv/memory" in labels:
            a = (a + str_to_u64(labels["scv/memory"])) & M64
    if node["total_sum"] < a:
        return 0
    return (((((node["total_sum"] - a) * 100) & M64) // node["total_sum"]) * 3) & M64


def score(mx, labels, node) -> int:
    """algorithm.go:28-39 + scheduler.go:128"""
    return u64_to_i64((basic_score(mx, labels, node) + allocate_score(node) + actual_score(node)) & M64)


def _wrap_i64(x: int) -> int:
    x &= M64
    return x - (1 << 64) if x >= I63 else x


def _go_div(a: int, b: int) -> int:
    q = abs(a) // abs(b)
    return _wrap_i64(-q if (a < 0) != (b < 0) else q)


def normalize(scores):
    """scheduler.go:132-157; returns (normalized list, lowest, highest) after the lowest-- step"""
    hi, lo = 0, scores[0]
    for s in scores:
        lo = min(lo, s)
        hi = max(hi, s)
    if hi == lo:
        lo = _wrap_i64(lo - 1)
    den = _wrap_i64(hi - lo)
    return [_go_div(_wrap_i64(_wrap_i64(s - lo) * 100), den) for s in scores], lo, hi


def cycle(labels, nodes):
    """Filter -> CollectMaxValues -> Score -> NormalizeScore -> lowest-index argmax."""
    N = len(nodes)
    feas = [yoda_filter(labels, nd) for nd in nodes]
    mx = collect_max(labels, nodes)
    raw = [score(mx, labels, nd) if nd["has_scv"] else 0 for nd in nodes]
    idx = [i for i in range(N) if feas[i]]
    norm = [0] * N
    lo = hi = 0
    chosen = -1
    if idx:
        nl, lo, hi = normalize([raw[i] for i in idx])
        best = -1
        for i, v in zip(idx, nl):
            norm[i] = v
            if v > best:
                best, chosen = v, i
    return {
        "feasible": feas, "raw": raw, "norm": norm,
        "max6": [mx["bandwidth"], mx["clock"], mx["core"], mx["free"], mx["power"], mx["total"]],
        "lowest": lo, "highest": hi, "count": len(idx), "chosen": chosen,
    }
