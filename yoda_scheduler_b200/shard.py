"""Node-sharded evaluation: host-side plumbing between the per-GPU contexts.

Each shard evaluates every pod against its own node range and produces four small per-pod partial
buffers (uint64, see csrc/common.cuh):

    max6[p][6]  collection.MaxValue of the shard            reduce: MAX   (after phase 0)
    hilo[p][2]  highest, ~lowest over the shard's feasible  reduce: MAX   (after phase 1)
    ckey[p]     2^32 - global index of the shard's choice   reduce: MAX   (after phase 2)
    count[p]    feasible nodes of the shard                 reduce: SUM   (after phase 2)

The encodings are order preserving as *unsigned* 64-bit integers, so an element-wise reduction
finishes them.  torch.distributed has no unsigned 64-bit reductions; `allreduce_u64` therefore flips
the sign bit around a signed MAX.  The same functions serve NCCL (device tensors) and gloo (CPU
tensors, used by the CPU tests of this logic).
"""
from __future__ import annotations

from typing import Optional

import numpy as np

XCHG_MAX, XCHG_SUM = 0, 1
SIGN = -(2**63)


def ckey_of(gidx):
    return (np.uint64(0xFFFFFFFF) - np.asarray(gidx, dtype=np.uint64)) + np.uint64(1)


def ckey_to_idx(k):
    k = np.asarray(k, dtype=np.uint64)
    idx = (np.uint64(0xFFFFFFFF) - (k - np.uint64(1))).astype(np.int64)
    return np.where(k == 0, -1, idx).astype(np.int32)


def encode_hilo(raw_feasible: np.ndarray):
    """highest (starts at 0, scheduler.go:134) and ~lowest of a shard's feasible scores; (0, 0) if none."""
    if len(raw_feasible) == 0:
        return np.uint64(0), np.uint64(0)
    hi = max(0, int(raw_feasible.max()))
    lo = int(raw_feasible.min())
    return np.uint64(hi), np.uint64(~lo & (2**64 - 1))


def decode_lo_hi(hi_u64, nlo_u64):
    """lowest/highest as scheduler.go:147-152 uses them, from the reduced pair."""
    if int(nlo_u64) == 0:
        return 0, 0
    hi = int(hi_u64)
    lo = ~int(nlo_u64) & (2**64 - 1)
    if lo >= 2**63:
        lo -= 2**64
    if hi == lo:
        lo -= 1
    return lo, hi


def allreduce_u64(t, op: int, group=None):
    """In-place element-wise MAX / SUM of a tensor holding uint64 bit patterns in int64 storage."""
    import torch.distributed as dist
    if op == XCHG_SUM:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return t
    t.bitwise_xor_(SIGN)           # unsigned order -> signed order
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    t.bitwise_xor_(SIGN)
    return t


class _DevBuf:
    """A raw device pointer as a CUDA array (so torch can wrap it without copying)."""

    def __init__(self, ptr: int, count: int):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<i8", "data": (ptr, False), "version": 3}


def device_tensor(ptr: int, count: int, device):
    import torch
    return torch.as_tensor(_DevBuf(ptr, count), device=device)


def init_nccl(ctx, rank: int, world: int, device):
    """Give the context its own NCCL communicator; the 128-byte id travels over torch.distributed."""
    import torch
    import torch.distributed as dist
    from .api import check
    idt = torch.zeros(128, dtype=torch.uint8, device=device)
    if rank == 0:
        buf = np.zeros(128, np.uint8)
        check(ctx.L.yoda_comm_unique_id(buf.ctypes.data))
        idt.copy_(torch.from_numpy(buf))
    dist.broadcast(idt, 0)
    buf = idt.cpu().numpy().copy()
    check(ctx.L.yoda_comm_init(ctx.h, buf.ctypes.data, rank, world))


def eval_host_driven(ctx, reqs, flags, device, group=None):
    """Sharded evaluation with the exchanges done by torch.distributed instead of in-library NCCL."""
    ctx.phase_begin(reqs, flags)
    for phase in (0, 1, 2):
        ctx.phase_run(phase)
        for ptr, count, op in ctx.phase_exchange(phase):
            allreduce_u64(device_tensor(ptr, count, device), op, group)
        import torch
        torch.cuda.synchronize(device)
    ctx.phase_run(3)
    return ctx.phase_finish()
