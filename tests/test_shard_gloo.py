"""Node sharding, host logic, on CPU: world_size-2 gloo processes run the shard protocol of
yoda_scheduler_b200/shard.py (encodings + all-reduces) with the oracle standing in for the kernels,
and the combined result must equal the unsharded oracle.  No GPU involved."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n_nodes, n_pods, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import oracle as ora
    from yoda_scheduler_b200 import shard, synth

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    per = n_nodes // world
    base = rank * per
    snap = synth.make_snapshot(per, node_offset=base)
    labels = synth.make_pod_labels(n_pods, 77)
    labels += [{"scv/memory": "999999"}, {"scv/number": "0"}, {}]        # nothing feasible / everything / no labels
    cl = ora.OracleCluster.from_snapshot(snap)
    P = len(labels)
    # phase 0: local maxima -> MAX
    local = cl.eval(labels, full=False)
    max6 = torch.from_numpy(local["max6"].view(np.int64).copy().reshape(-1))
    shard.allreduce_u64(max6, shard.XCHG_MAX)
    max6 = max6.numpy().view(np.uint64).reshape(P, 6)
    # phase 1: score under the global maxima, local highest / ~lowest -> MAX
    feas, raw = [], []
    hilo = np.zeros((P, 2), np.uint64)
    for p, lab in enumerate(labels):
        f, r = cl.score_with_max(lab, max6[p])
        feas.append(f.astype(bool)); raw.append(r)
        hilo[p] = shard.encode_hilo(r[f.astype(bool)])
    t = torch.from_numpy(hilo.view(np.int64).copy().reshape(-1))
    shard.allreduce_u64(t, shard.XCHG_MAX)
    hilo = t.numpy().view(np.uint64).reshape(P, 2)
    # phase 2: local choice (lowest index holding the global top score) -> MAX, counts -> SUM
    ckey = np.zeros(P, np.uint64)
    count = np.zeros(P, np.uint64)
    lows, highs = [], []
    for p in range(P):
        lo, hi = shard.decode_lo_hi(hilo[p, 0], hilo[p, 1])
        lows.append(lo); highs.append(hi)
        top = np.nonzero(feas[p] & (raw[p] == hi))[0]
        if len(top) and hilo[p, 1] != 0:
            ckey[p] = shard.ckey_of(base + top[0])
        count[p] = feas[p].sum()
    t = torch.from_numpy(ckey.view(np.int64).copy())
    shard.allreduce_u64(t, shard.XCHG_MAX)
    chosen = shard.ckey_to_idx(t.numpy().view(np.uint64))
    t = torch.from_numpy(count.view(np.int64).copy())
    shard.allreduce_u64(t, shard.XCHG_SUM)
    count = t.numpy().view(np.uint64)
    if rank == 0:
        np.savez(out_path, max6=max6, lowest=np.array(lows, np.int64), highest=np.array(highs, np.int64), chosen=chosen, count=count)
    dist.barrier()
    dist.destroy_process_group()


def test_two_shards_equal_single(tmp_path):
    import torch.multiprocessing as mp
    from oracle import oracle as ora
    from yoda_scheduler_b200 import synth

    n_nodes, n_pods, world = 600, 60, 2
    port = 29500 + os.getpid() % 2000
    out = str(tmp_path / "sharded.npz")
    mp.spawn(_worker, args=(world, port, n_nodes, n_pods, out), nprocs=world, join=True)
    got = np.load(out)
    labels = synth.make_pod_labels(n_pods, 77) + [{"scv/memory": "999999"}, {"scv/number": "0"}, {}]
    want = ora.OracleCluster.from_snapshot(synth.make_snapshot(n_nodes)).eval(labels, full=False)
    assert np.array_equal(got["max6"], want["max6"])
    assert np.array_equal(got["lowest"], want["lowest"])
    assert np.array_equal(got["highest"], want["highest"])
    assert np.array_equal(got["chosen"], want["chosen"])
    assert np.array_equal(got["count"], want["count"])


def test_encodings():
    from yoda_scheduler_b200 import shard
    assert shard.ckey_to_idx(shard.ckey_of([0, 5, 123456])).tolist() == [0, 5, 123456]
    assert shard.ckey_to_idx(np.zeros(2, np.uint64)).tolist() == [-1, -1]
    assert shard.ckey_of(3) > shard.ckey_of(4)                  # MAX picks the lowest index
    hi, nlo = shard.encode_hilo(np.array([7, 7], np.int64))
    assert shard.decode_lo_hi(hi, nlo) == (6, 7)                 # all equal -> lowest-- (scheduler.go:147-149)
    hi, nlo = shard.encode_hilo(np.array([0, 0], np.int64))
    assert shard.decode_lo_hi(hi, nlo) == (-1, 0)
    assert shard.decode_lo_hi(*shard.encode_hilo(np.array([], np.int64))) == (0, 0)
    assert shard.decode_lo_hi(*shard.encode_hilo(np.array([3, 10], np.int64))) == (3, 10)
