"""Node sharding on the GPU: shards driven through the phase API must reproduce the single-context
result bit for bit (and hence the oracle)."""
import os
import sys

import numpy as np
import pytest

from oracle import oracle as ora
from parity import assert_matches_oracle
from yoda_scheduler_b200 import api, shard, synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("wide", [None, "stat", "cardnum"])
@pytest.mark.parametrize("flags", [api.EVAL_FULL, api.EVAL_FULL | api.EVAL_FORCE_DIRECT, api.EVAL_DECIDE])
@pytest.mark.parametrize("n_shards", [2, 3])
def test_shards_on_one_gpu_equal_single(n_shards, flags, wide):
    """All shards live on cuda:0 and run one after the other; the exchange between phases is an
    element-wise max / sum over the shards' partial buffers (what NCCL does across GPUs).
    wide: one node of the first shard has a static score beyond 2^32 ("stat": it becomes `highest` of most pods,
    so the other, all-narrow shards must rescale their rows with 64-bit arithmetic) or a 40-bit CardNumber."""
    import torch
    dev = torch.device("cuda", 0)
    N, P = 3000, 500
    snap = synth.make_snapshot(N)
    if wide:
        node = int(np.nonzero((snap.has_scv == 1) & (snap.n_cards == 8) & (snap.card_number == 8))[0][3])
        if wide == "stat":
            snap.free_memory_sum[node] = np.uint64(1) << np.uint64(50)
        else:
            snap.card_number[node] = np.uint64(1) << np.uint64(40)
    reqs = synth.make_requests(P, 31)
    labels = synth.make_pod_labels(P, 31)
    with api.Context(N) as single:
        single.load_snapshot(snap)
        want = single.eval(reqs, flags)
        want_m = single.fetch_all(P) if flags & api.EVAL_FULL else None
    bounds = np.linspace(0, N, n_shards + 1).astype(int)
    ctxs = []
    for r in range(n_shards):
        lo, hi = int(bounds[r]), int(bounds[r + 1])
        c = api.Context(hi - lo, node_base=lo, shard_rank=r, shard_count=n_shards)
        c.load_snapshot(snap.slice(lo, hi))
        c.phase_begin(reqs, flags)
        ctxs.append(c)
    for phase in (0, 1, 2):
        for c in ctxs:
            c.phase_run(phase)
        bufs = [c.phase_exchange(phase) for c in ctxs]
        for k in range(len(bufs[0])):
            ts = [shard.device_tensor(b[k][0], b[k][1], dev) for b in bufs]
            op = bufs[0][k][2]
            if op == shard.XCHG_SUM:
                red = torch.stack(ts).sum(0)
            else:
                red = (torch.stack(ts) ^ shard.SIGN).max(0).values ^ shard.SIGN   # unsigned max
            for t in ts:
                t.copy_(red)
        torch.cuda.synchronize()
    got = []
    for c in ctxs:
        c.phase_run(3)
        got.append(c.phase_finish())
    for g in got:
        assert np.array_equal(g, want)          # every shard ends with the same global per-pod results
    if flags & api.EVAL_FULL:
        mats = [c.fetch_all(P) for c in ctxs]
        for k in ("feasible", "raw", "norm"):
            assert np.array_equal(np.concatenate([m[k] for m in mats], axis=1), want_m[k]), k
        assert_matches_oracle(snap, labels, want, want_m)
    for c in ctxs:
        c.close()


def _widen(snap, lo):
    """Give the first node at or after `lo` that has an SCV a CardNumber beyond 32 bits: its shard leaves the narrow
    fast path, the other shards do not (a wide card field would travel to them through the maxima)."""
    node = lo + int(np.nonzero(snap.has_scv[lo:] == 1)[0][0])
    snap.card_number[node] = np.uint64(1) << np.uint64(40)


def _nccl_worker(rank, world, port, N, P, out_dir):
    try:
        _nccl_worker_body(rank, world, port, N, P, out_dir)
    except BaseException:   # a rank that dies quietly leaves the other one waiting in a collective for minutes
        import traceback
        traceback.print_exc()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(1)


def _nccl_worker_body(rank, world, port, N, P, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from yoda_scheduler_b200 import api, shard, synth
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world, device_id=dev)
    per = N // world
    ctx = api.Context(per, device=rank, node_base=rank * per, shard_rank=rank, shard_count=world)
    shard.init_nccl(ctx, rank, world, dev)
    snap = synth.make_snapshot(per, node_offset=rank * per)
    ctx.load_snapshot(snap)
    reqs = synth.make_requests(P, 31)
    res = ctx.eval(reqs, api.EVAL_FULL)                       # exchanges run inside the library (NCCL)
    res2 = shard.eval_host_driven(ctx, reqs, api.EVAL_FULL, dev)   # same through torch.distributed
    assert np.array_equal(res, res2)
    np.save(os.path.join(out_dir, f"res{rank}.npy"), res)
    # DECIDE: every shard is on the table path, so they agree on the one-pass variant (choices out of the score pass)
    dec = ctx.eval(reqs, api.EVAL_DECIDE)
    assert ctx.timing().path == 1
    np.save(os.path.join(out_dir, f"dec{rank}.npy"), dec)
    # DECIDE with one shard on the exact kernels: the shards must agree on the two-pass protocol instead
    if rank == 1:
        _widen(snap, 0)
        ctx.load_snapshot(snap)
    mixed = ctx.eval(reqs, api.EVAL_DECIDE)
    assert ctx.timing().path == 1      # a wide node no longer costs its shard the table path
    np.save(os.path.join(out_dir, f"mix{rank}.npy"), mixed)
    mixed_full = ctx.eval(reqs, api.EVAL_FULL)
    assert np.array_equal(mixed_full, mixed)
    # FULL with a class table that does not fit the packed format on ONE shard (a 4e9 clock: Clock*100/MaxBandwidth leaves
    # 21 bits): that shard's kernels switch to the exact path on the device (no host round trip), the other shard's do not
    snap = synth.make_snapshot(per, node_offset=rank * per)
    if rank == 1:
        node = int(np.nonzero((snap.has_scv == 1) & (snap.n_cards == 8))[0][5])
        snap.clock[node, :] = np.uint64(4_000_000_000)
    ctx.load_snapshot(snap)
    over = ctx.eval(reqs, api.EVAL_FULL)
    assert ctx.timing().path == (0 if rank == 1 else 1), ctx.timing().path
    np.save(os.path.join(out_dir, f"over{rank}.npy"), over)
    np.save(os.path.join(out_dir, f"over_raw{rank}.npy"), np.stack([ctx.fetch_row(p, api.ROW_RAW, per) for p in range(40)]))
    np.save(os.path.join(out_dir, f"over_norm{rank}.npy"), np.stack([ctx.fetch_row(p, api.ROW_NORM, per) for p in range(40)]))
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("xchg", ["nccl", "peer"])
def test_two_gpus_nccl(tmp_path, xchg, monkeypatch):
    """Two processes, two GPUs, the exchanges inside the library: over NCCL (default) and over peer memory
    (YODA_B200_XCHG=peer: CUDA-IPC mapped arenas, publish / reduce kernels of csrc/xchg_kernels.cuh)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    import torch.multiprocessing as mp
    N, P, world = 4000, 600, 2
    port = 29500 + (os.getpid() + (7 if xchg == "peer" else 0)) % 2000
    monkeypatch.setenv("YODA_B200_XCHG", xchg)     # the spawned shards inherit it
    mp.spawn(_nccl_worker, args=(world, port, N, P, str(tmp_path)), nprocs=world, join=True)
    reqs = synth.make_requests(P, 31)
    snap = synth.make_snapshot(N)
    with api.Context(N) as single:
        single.load_snapshot(snap)
        want = single.eval(reqs, api.EVAL_FULL)
        want_dec = single.eval(reqs, api.EVAL_DECIDE)
        assert np.array_equal(want, want_dec)
        _widen(snap, N // world)
        single.load_snapshot(snap)
        want_mix = single.eval(reqs, api.EVAL_DECIDE)
        assert single.timing().path == 1
        snap = synth.make_snapshot(N)
        node = N // world + int(np.nonzero((snap.has_scv[N // world:] == 1) & (snap.n_cards[N // world:] == 8))[0][5])
        snap.clock[node, :] = np.uint64(4_000_000_000)
        single.load_snapshot(snap)
        want_over = single.eval(reqs, api.EVAL_FULL | api.EVAL_FORCE_DIRECT)
        want_raw = np.stack([single.fetch_row(p, api.ROW_RAW, N) for p in range(40)])
        want_norm = np.stack([single.fetch_row(p, api.ROW_NORM, N) for p in range(40)])
    o = ora.OracleCluster.from_snapshot(snap).eval(synth.make_pod_labels(P, 31), threads=8, full=False)
    assert np.array_equal(want_over["chosen"], o["chosen"]) and np.array_equal(want_over["highest"], o["highest"])
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"res{r}.npy"), want)
        assert np.array_equal(np.load(tmp_path / f"dec{r}.npy"), want)
        assert np.array_equal(np.load(tmp_path / f"mix{r}.npy"), want_mix)
        assert np.array_equal(np.load(tmp_path / f"over{r}.npy"), want_over)
    assert np.array_equal(np.concatenate([np.load(tmp_path / f"over_raw{r}.npy") for r in range(world)], axis=1), want_raw)
    assert np.array_equal(np.concatenate([np.load(tmp_path / f"over_norm{r}.npy") for r in range(world)], axis=1), want_norm)


def test_pod_sharding_equals_single():
    """The other way to use several GPUs (bench.py --shard pods): the snapshot is
    replicated, every GPU takes its own slice of the pending pods, nothing is exchanged.  Two contexts
    on cuda:0 stand in for two GPUs: their results, concatenated, are the single-batch results."""
    N, P = 3000, 601
    snap = synth.make_snapshot(N)
    reqs = synth.make_requests(P, 77)
    with api.Context(N) as single:
        single.load_snapshot(snap)
        want = single.eval(reqs, api.EVAL_FULL)
        want_m = single.fetch_all(P)
    cut = 300
    got, rows = [], {"feasible": [], "raw": [], "norm": []}
    for part in (reqs[:cut], reqs[cut:]):
        with api.Context(N) as c:
            c.load_snapshot(snap)
            got.append(c.eval(np.ascontiguousarray(part), api.EVAL_FULL))
            m = c.fetch_all(len(part))
            for k in rows:
                rows[k].append(m[k])
    assert np.array_equal(np.concatenate(got), want)
    for k in rows:
        assert np.array_equal(np.concatenate(rows[k]), want_m[k]), k
    assert_matches_oracle(snap, synth.make_pod_labels(P, 77), want, want_m, what="pod-sharded")
