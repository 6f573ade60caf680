"""Host-side label handling (the scv/* label contract).

filter.go:60-82 turns a label string into an integer with strconv.Atoi and maps any error
to 0.  Parsing stays on the host, once per pod (the reference parses once per pod x node).
"""
from __future__ import annotations

from typing import Mapping, Optional, Sequence, Tuple

import numpy as np

from .snapshot import POD_REQ

U64 = (1 << 64) - 1
LABEL_NUMBER = "scv/number"      # filter.go:12
LABEL_MEMORY = "scv/memory"      # filter.go:19, algorithm.go:77
LABEL_CLOCK = "scv/clock"        # filter.go:36
LABEL_PRIORITY = "scv/priority"  # sort.go:13

HAS_NUMBER, HAS_MEMORY, HAS_CLOCK = 1, 2, 4


def go_atoi(s: str) -> Tuple[int, Optional[str]]:
    """[3P] Go strconv.Atoi: (value, error) with error in {None, "syntax", "range"}.

    On a range error Go returns the clamped value; on a syntax error 0.
    """
    if s == "":
        return 0, "syntax"
    neg = s[0] == "-"
    body = s[1:] if s[0] in "+-" else s
    if body == "":
        return 0, "syntax"
    v = 0
    for ch in body:
        if not ("0" <= ch <= "9"):
            return 0, "syntax"
        v = v * 10 + (ord(ch) - 48)
    if neg:
        v = -v
    if v > (1 << 63) - 1:
        return (1 << 63) - 1, "range"
    if v < -(1 << 63):
        return -(1 << 63), "range"
    return v, None


def str_to_uint64(s: str) -> int:
    """filter.StrToUint64 / strToUint (filter.go:60-74)."""
    v, err = go_atoi(s)
    return 0 if err else v & U64


def uint64_to_int64(u: int) -> int:
    """filter.Uint64ToInt64 (filter.go:84-86): >= 2^63 is a range error -> 0."""
    u &= U64
    return 0 if u >> 63 else u


def pod_priority(labels: Mapping[str, str]) -> int:
    """sort.GetPodPriority (sort.go:12-18): the error is dropped, the clamped value is kept."""
    if LABEL_PRIORITY in labels:
        return go_atoi(labels[LABEL_PRIORITY])[0]
    return 0


def pods_to_requests(pods: Sequence[Mapping[str, str]]) -> np.ndarray:
    """Label maps -> yoda_pod_req array (the requirement vectors uploaded to the device)."""
    out = np.zeros(len(pods), dtype=POD_REQ)
    for i, lab in enumerate(pods):
        f = 0
        if LABEL_NUMBER in lab:
            f |= HAS_NUMBER
            out["number"][i] = str_to_uint64(lab[LABEL_NUMBER])
        if LABEL_MEMORY in lab:
            f |= HAS_MEMORY
            out["memory"][i] = str_to_uint64(lab[LABEL_MEMORY])
        if LABEL_CLOCK in lab:
            f |= HAS_CLOCK
            out["clock"][i] = str_to_uint64(lab[LABEL_CLOCK])
        out["flags"][i] = f
    return out
