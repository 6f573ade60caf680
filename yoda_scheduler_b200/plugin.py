"""Python mirror of the plugin surface of pkg/yoda/scheduler.go on top of the C ABI.

Same method names, argument meaning, status codes and messages as the Go plugin (and as the C++
mirror in host/yoda_plugin.cpp, which is what the benchmark harness drives); used by the tests so
that they read like tests of the reference plugin.  Citations: pkg/yoda/scheduler.go.
"""
from __future__ import annotations

import threading
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import api, labels

NAME = "yoda"                      # scheduler.go:23-25
SUCCESS, ERROR, UNSCHEDULABLE = 0, 1, 2
MAX_NODE_SCORE = 100               # framework.MaxNodeScore
_MEMO = "yoda/b200"


@dataclass
class Status:
    code: int = SUCCESS
    message: str = ""

    def is_success(self) -> bool:
        return self.code == SUCCESS


@dataclass
class Pod:
    name: str = ""
    namespace: str = "default"
    labels: Dict[str, str] = field(default_factory=dict)


@dataclass
class NodeInfo:
    name: str
    pods: List[Pod] = field(default_factory=list)


@dataclass
class NodeScore:
    name: str
    score: int = 0


class CycleState(dict):
    """framework.CycleState: a per-cycle key/value store with its own lock."""

    def __init__(self):
        super().__init__()
        self.lock = threading.Lock()


@dataclass
class MaxValue:                    # collection.go:14-21
    MaxBandwidth: int = 1
    MaxClock: int = 1
    MaxCore: int = 1
    MaxFreeMemory: int = 1
    MaxPower: int = 1
    MaxTotalMemory: int = 1


class Handle:
    """framework.Handle reduced to SnapshotSharedLister().NodeInfos().Get (scheduler.go:111)."""

    def __init__(self):
        self.node_infos: Dict[str, NodeInfo] = {}

    def get(self, name: str) -> Optional[NodeInfo]:
        return self.node_infos.get(name)


class Yoda:
    def __init__(self, handle: Handle, max_nodes: int = 1 << 16, device: int = 0):
        """scheduler.go:46-74 New: here the "cache" is the device snapshot fed by on_scv_* / on_pod_*."""
        self.handle = handle
        self.ctx = api.Context(max_nodes, device=device)
        self.index: Dict[str, int] = {}
        self.names: List[str] = []
        self.has_scv: List[bool] = []
        self.alloc: List[int] = []
        self.dirty = False

    def close(self):
        self.ctx.close()

    def name(self) -> str:         # scheduler.go:42-44
        return NAME

    # ---- informer events ---------------------------------------------------------------------
    def _idx(self, name: str) -> int:
        if name not in self.index:
            self.index[name] = len(self.names)
            self.names.append(name)
            self.has_scv.append(False)
            self.alloc.append(0)
        return self.index[name]

    def on_scv_upsert(self, name: str, status: dict):
        """status: {"CardNumber", "FreeMemorySum", "TotalMemorySum", "CardList": [{"Health", "FreeMemory", ...}]}"""
        i = self._idx(name)
        rec = np.zeros(1, api.NODE_REC)
        cards = status["CardList"]
        rec["card_number"], rec["free_memory_sum"], rec["total_memory_sum"] = status["CardNumber"], status["FreeMemorySum"], status["TotalMemorySum"]
        rec["n_cards"] = len(cards)
        hm = 0
        for g, c in enumerate(cards[:8]):
            hm |= (1 << g) if c["Health"] == "Healthy" else 0   # filter.go:53,57
            for src, dst in (("FreeMemory", "free_memory"), ("TotalMemory", "total_memory"), ("Clock", "clock"), ("Core", "core"),
                             ("Bandwidth", "bandwidth"), ("Power", "power")):
                rec[dst][0, g] = c[src]
        rec["healthy_mask"] = hm
        try:
            self.ctx.upsert(i, rec)
        except api.YodaError:
            # refused (more than 8 cards; TotalMemorySum == 0, where the reference would panic): the node must not keep
            # answering from its previous record
            self.ctx.remove(i)
            self.has_scv[i] = False
            self.dirty = True
            raise
        self.has_scv[i] = True
        self.dirty = True

    def on_scv_delete(self, name: str):
        i = self._idx(name)
        self.ctx.remove(i)
        self.has_scv[i] = False
        self.dirty = True

    def on_pod_add(self, node: str, pod: Pod, sign: int = 1):
        if labels.LABEL_MEMORY not in pod.labels:               # algorithm.go:77
            return
        i = self._idx(node)
        self.alloc[i] = (self.alloc[i] + sign * labels.str_to_uint64(pod.labels[labels.LABEL_MEMORY])) & labels.U64
        self.ctx.set_alloc(i, self.alloc[i])
        self.dirty = True

    def on_pod_delete(self, node: str, pod: Pod):
        self.on_pod_add(node, pod, -1)

    def sync(self):
        if self.dirty:
            self.ctx.commit()
            self.dirty = False

    # ---- one device pass per cycle, memoised in CycleState ------------------------------------
    def _memo(self, state: CycleState, pod: Pod):
        with state.lock:                                        # as collection.go:53-55 does for "Max"
            if _MEMO not in state:
                self.sync()
                res = self.ctx.eval(labels.pods_to_requests([pod.labels]), api.EVAL_FULL)
                n = self.ctx.node_count()
                n = min(n, len(self.has_scv))
                # the cycle's frozen view: rows plus has_scv as of the evaluation; a node registered later is not part of it
                state[_MEMO] = (res[0], self.ctx.fetch_row(0, api.ROW_FEASIBLE_BITS, n), self.ctx.fetch_row(0, api.ROW_RAW, n),
                                self.ctx.fetch_row(0, api.ROW_NORM, n), list(self.has_scv[:n]))
            return state[_MEMO]

    @staticmethod
    def _known(memo, i) -> bool:
        return i is not None and i < len(memo[4]) and memo[4][i]

    # ---- plugin methods -----------------------------------------------------------------------
    def filter(self, state: CycleState, pod: Pod, node: NodeInfo) -> Status:
        """scheduler.go:76-93"""
        memo = self._memo(state, pod)
        bits = memo[1]
        i = self.index.get(node.name)
        if not self._known(memo, i):
            return Status(UNSCHEDULABLE, f'Node:{node.name} Scv.core.run-linux.com "{node.name}" not found')   # :83
        if i < len(bits) and bits[i]:
            return Status(SUCCESS, "")
        return Status(UNSCHEDULABLE, "Node:" + node.name)      # :92

    def post_filter(self, state: CycleState, pod: Pod) -> Status:
        """scheduler.go:95-103 + collection.go:30-57: publishes the maxima under "Max"."""
        r = self._memo(state, pod)[0]
        with state.lock:
            state["Max"] = MaxValue(int(r["max_bandwidth"]), int(r["max_clock"]), int(r["max_core"]), int(r["max_free_memory"]),
                                    int(r["max_power"]), int(r["max_total_memory"]))
        return Status(SUCCESS, "")

    def less(self, a: Pod, b: Pod) -> bool:
        """scheduler.go:105-107 / sort.go:8-18"""
        return labels.pod_priority(a.labels) > labels.pod_priority(b.labels)

    def score(self, state: CycleState, pod: Pod, node_name: str) -> Tuple[int, Status]:
        """scheduler.go:109-130"""
        if self.handle.get(node_name) is None:
            return 0, Status(ERROR, f'getting node "{node_name}" from Snapshot: nodeinfo not found for node name "{node_name}"')
        memo = self._memo(state, pod)
        i = self.index.get(node_name)
        if not self._known(memo, i):                            # :117-122; also a node registered after the evaluation
            return 0, Status(ERROR, f'Score Node Error: Scv.core.run-linux.com "{node_name}" not found')
        if "Max" not in state:                                  # algorithm.go:29-33
            return 0, Status(ERROR, "Score Node Error: not found")
        return int(memo[2][i]), Status(SUCCESS, "")

    def normalize_score(self, state: CycleState, pod: Pod, scores: List[NodeScore]) -> Optional[Status]:
        """scheduler.go:132-157: in place, returns None (nil); an empty list raises like the Go index panic."""
        if not scores:
            raise IndexError("index out of range [0] with length 0")
        memo = state.get(_MEMO)
        own = memo is not None and len(scores) == int(memo[0]["feasible"])
        if own:
            for s in scores:
                i = self.index.get(s.name)
                if i is None or i >= len(memo[1]) or not memo[1][i] or int(memo[2][i]) != s.score or int(memo[3][i]) == 255:
                    own = False
                    break
        if own:
            for s in scores:
                s.score = int(memo[3][self.index[s.name]])
            return None
        vals = self.ctx.normalize_scores([s.score for s in scores])   # the framework's list is not Yoda's own feasible set
        for s, v in zip(scores, vals):
            s.score = int(v)
        return None

    def score_extensions(self):
        return self
