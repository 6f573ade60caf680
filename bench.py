#!/usr/bin/env python
"""bench.py -- pod x node Filter+Score evaluations per second on the BASELINE.json workload.

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # CPU arm: the oracle port of the Go plugin
    torchrun ... bench.py --gpus N ...                       # node-sharded, one rank per GPU, NCCL

One "step" is one scheduling pass of the whole pending batch over the whole snapshot:
Filter (every node) -> CollectMaxValues -> Score -> NormalizeScore -> node choice, in FULL mode
(feasible bits + int64 scores + uint8 normalized scores materialised in HBM).
N=1 workload = BASELINE config 3 (50k nodes x 8 cards, 100k pods).  N>1 = the same per-GPU shard
size (50k nodes per GPU, node-sharded; every rank sees all pods) -> weak scaling.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pod_x_node_filter_score_evals_per_sec"
UNIT = "evals/s"
W_BYTES = 9.125      # int64 raw + uint8 normalized + 1 feasible bit per (pod, node)
R_NODE, R_POD, R_OUT = 208, 32, 72


W_BYTES_RAW32 = 5.125  # YODA_EVAL_RAW32: the same scores as uint32 (SURVEY.md 8(d) allows restating W this way)


def algorithmic_bytes(P, N, w=W_BYTES):
    """N*R_node + P*R_pod + P*R_out + P*N*W."""
    return N * R_NODE + P * R_POD + P * R_OUT + P * N * w


def profiled_traffic(config, mode, path):
    """dram bytes of the output pass per step from the committed ncu --set full capture (same workload)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        if t["config"] == f"{config} {mode}" and path == 1:
            return float(t["dram_bytes_read"]) + float(t["dram_bytes_write"])
    except Exception:
        pass
    return None


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index), "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def workload(args, rank, world):
    from yoda_scheduler_b200 import synth
    n_nodes, n_pods, pod_seed = synth.CONFIGS[args.config]
    if args.config == "c4":
        n_nodes //= world   # config 4 is ONE 1M-node cluster node-sharded over the GPUs (strong scaling)
    if args.nodes:
        n_nodes = args.nodes
    if args.pods:
        n_pods = args.pods
    return n_nodes, n_pods, pod_seed


def cpu_arm(snap, labels_fn, n_nodes, threads, target_s, faithful=True, cl=None):
    """Time the oracle (CPU restatement of the Go plugin) on a bounded pod sample x all nodes."""
    from oracle import oracle as ora
    cl = cl or ora.OracleCluster.from_snapshot(snap)
    probe = labels_fn(max(threads, 8))
    t0 = time.perf_counter()
    cl.eval(probe, faithful=faithful, threads=threads, full=False)
    dt = time.perf_counter() - t0
    per_pod = dt / len(probe)
    n = int(max(len(probe), min(target_s / max(per_pod, 1e-9), 20000)))
    return cl, n, per_pod


def run_reference(args):
    """--impl reference: the reference's own CPU algorithm (oracle port; the Go plugin itself cannot be
    built in this image), all host threads, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from yoda_scheduler_b200 import synth
    n_nodes, n_pods, pod_seed = workload(args, 0, 1)
    cores = os.cpu_count() or 1
    snap = synth.make_snapshot(n_nodes)
    all_labels = synth.make_pod_labels(min(n_pods, 4096), pod_seed)
    labels_fn = lambda k: [all_labels[i % len(all_labels)] for i in range(k)]
    budget = 150.0 / max(args.steps + args.warmup, 1)
    cl, sample, _ = cpu_arm(snap, labels_fn, n_nodes, cores, min(budget, 10.0))
    labels = labels_fn(sample)
    for _ in range(args.warmup):
        cl.eval(labels, faithful=True, threads=cores, full=False)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cl.eval(labels, faithful=True, threads=cores, full=False)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    value = sample * n_nodes / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": f"{args.config}: {n_nodes} nodes x 8 cards, {n_pods} pods; each step a {sample}-pod sample x all nodes"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample} pods x {n_nodes} nodes per step, faithful driver (per-call Scv copies, label lookup + Atoi, string health compare; logging excluded)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def sharded_parity_check(ctx, api, synth, dist, torch, dev, rank, world, n_nodes, reqs_np, flags, n_check=256, n_rows=8):
    """N>1 self-check: the first 256 pods of the batch, evaluated node-sharded (what the timed region ran) and on ONE
    unsharded context holding all N x n_nodes nodes (rank 0's GPU), must give identical per-pod results; rank 0 also
    compares its shard's slice of the first 8 pods' rows.  Makes the driver's scaling runs correctness evidence too."""
    import numpy as np
    sub = np.ascontiguousarray(reqs_np[:n_check])
    got = ctx.eval(sub, flags)                       # collective: every rank takes part
    rows = [(ctx.fetch_row(p, api.ROW_FEASIBLE_BITS, n_nodes), ctx.fetch_row(p, api.ROW_RAW, n_nodes), ctx.fetch_row(p, api.ROW_NORM, n_nodes))
            for p in range(min(n_rows, len(sub)))] if (flags & api.EVAL_FULL) else []
    ok = torch.ones(1, device=dev, dtype=torch.int32)
    info = {"pods": int(len(sub)), "rows": len(rows), "against": f"one unsharded context of {n_nodes * world} nodes on rank 0"}
    if rank == 0:
        whole = synth.make_snapshot(n_nodes * world)
        with api.Context(n_nodes * world, device=dev.index) as single:
            single.load_snapshot(whole)
            want = single.eval(sub, flags & ~api.EVAL_RAW32)
            equal = bool(np.array_equal(got, want))
            for p, (b, r, nm) in enumerate(rows):
                equal = equal and np.array_equal(b, single.fetch_row(p, api.ROW_FEASIBLE_BITS, n_nodes * world)[:n_nodes]) \
                    and np.array_equal(r, single.fetch_row(p, api.ROW_RAW, n_nodes * world)[:n_nodes]) \
                    and np.array_equal(nm, single.fetch_row(p, api.ROW_NORM, n_nodes * world)[:n_nodes])
        ok[0] = 1 if equal else 0
    dist.broadcast(ok, 0)
    info["equal"] = bool(ok.item())
    return info


def fallback_paths(ctx, api, synth, torch, stream, d_reqs, snap, n_nodes, n_pods, reqs_np, flags, headline_ms):
    """What the headline's dtype hides: (1) the exact-u64 kernels on a 2000-pod slice of the same batch (the path a batch
    takes when its class maxima leave 32 bits), extrapolated per evaluation; (2) the same batch with ONE node carrying a
    40-bit CardNumber -- the table path keeps every other node and the exact kernels take just that one."""
    import numpy as np
    out = {}
    sl = 2000
    for _ in range(2):
        ctx.eval_device(d_reqs.data_ptr(), sl, flags | api.EVAL_FORCE_DIRECT)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(3):
            ctx.eval_device(d_reqs.data_ptr(), sl, flags | api.EVAL_FORCE_DIRECT)
        e1.record(stream)
    torch.cuda.synchronize()
    dms = e0.elapsed_time(e1) / 3
    out["direct_path"] = {"pods": sl, "ms_per_step": dms, "value": sl * n_nodes / (dms * 1e-3), "unit": UNIT,
                          "note": "YODA_EVAL_FORCE_DIRECT on a 2000-pod slice: exact-u64 kernels with real 64-bit divisions, the whole-batch fallback"}
    node = int(np.nonzero((snap.has_scv == 1) & (snap.card_number == 8))[0][7])
    rec = snap.records(np.array([node]))
    keep = rec["card_number"].copy()
    rec["card_number"] = np.uint64(1) << np.uint64(40)
    ctx.upsert_many(np.array([node], np.uint32), rec)
    ctx.commit()
    for _ in range(3):
        ctx.eval_device(d_reqs.data_ptr(), n_pods, flags)
    ctx.sync()
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(10):
            ctx.eval_device(d_reqs.data_ptr(), n_pods, flags)
        e1.record(stream)
    torch.cuda.synchronize()
    wms = e0.elapsed_time(e1) / 10
    out["one_wide_node"] = {"ms_per_step": wms, "value": n_pods * n_nodes / (wms * 1e-3), "unit": UNIT, "path": int(ctx.timing().path),
                            "vs_headline": headline_ms / wms, "note": f"node {node} with CardNumber = 2^40: table path for the other nodes, exact kernels for this one"}
    rec["card_number"] = keep
    ctx.upsert_many(np.array([node], np.uint32), rec)
    ctx.commit()
    return out


def c4_strong(args, api, synth, dist, torch, dev, rank, world, local_rank):
    """BASELINE config 4 beside the weak-scaled headline: ONE 1M-node cluster node-sharded over the N GPUs (1M / N nodes
    each), 10k pods, FULL mode -- the strong-scaling curve north_star asks for.  Timed like the headline."""
    import numpy as np
    total_nodes, n_pods, pod_seed = synth.CONFIGS["c4"]
    per = total_nodes // world
    base = rank * per
    ctx = api.Context(per, device=local_rank, node_base=base, shard_rank=rank, shard_count=world)
    if world > 1:
        from yoda_scheduler_b200 import shard
        shard.init_nccl(ctx, rank, world, dev)
    ctx.load_snapshot(synth.make_snapshot(per, node_offset=base))
    reqs = synth.make_requests(n_pods, pod_seed)
    d_reqs = torch.from_numpy(reqs.view(np.uint8).copy()).to(dev)
    stream = torch.cuda.Stream(device=dev)
    api.check(ctx.L.yoda_set_stream(ctx.h, stream.cuda_stream))
    steps = max(3, min(args.steps, 10))
    for _ in range(3):
        ctx.eval_device(d_reqs.data_ptr(), n_pods, api.EVAL_FULL)
    ctx.sync()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    phase = {"prep": 0.0, "max": 0.0, "score": 0.0, "write": 0.0, "total": 0.0}
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(steps):
            ctx.eval_device(d_reqs.data_ptr(), n_pods, api.EVAL_FULL)
            t = ctx.timing()
            for k in phase:
                phase[k] += getattr(t, k + "_ms")
        e1.record(stream)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    path = int(ctx.timing().path)
    ctx.close()
    if world > 1:
        tt = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    peak, _ = measured_peak()
    step_bytes = algorithmic_bytes(n_pods, per)        # per GPU: every GPU writes its own slice of the matrices
    return {"workload": f"c4: {total_nodes} nodes x 8 cards node-sharded over {world} GPU(s) ({per} per GPU), {n_pods} pods, mode=full",
            "scaling": "strong", "steps": steps, "ms_per_step": ms, "value": float(n_pods) * total_nodes / (ms * 1e-3), "unit": UNIT,
            "phases_ms": {k: v / steps for k, v in phase.items()}, "step_achieved_per_gpu": step_bytes / (ms * 1e-3) / 1e9,
            "step_frac": step_bytes / (ms * 1e-3) / 1e9 / peak, "path": "narrow-fast" if path == 1 else "direct-u64"}


def plugin_cycle(n_nodes, cycles=64):
    """One pod through the plugin surface the way kube-scheduler calls it (Filter on every node from 16 threads ->
    PostFilter -> Score on the feasible nodes -> NormalizeScore), C++ mirror of pkg/yoda/scheduler.go over the C ABI,
    rows fetched to the host included: p50/p99 of the whole cycle (yoda_scheduler_b200/host/harness.cpp)."""
    try:
        from yoda_scheduler_b200 import _build
        exe = _build.build_host()
        r = subprocess.run([exe, "--nodes", str(n_nodes), "--pods", str(cycles), "--cycles", str(cycles), "--threads", "16"],
                           capture_output=True, text=True, timeout=600)
        info = json.loads(r.stdout.strip().splitlines()[-1])
        return {"nodes": n_nodes, "cycles": int(info["cycles_checked"]), "threads": int(info["threads"]), "p50_ms": info["framework_cycle_p50_ms"],
                "p99_ms": info["framework_cycle_p99_ms"], "mean_ms": info["framework_cycle_ms"], "mismatches": int(info["mismatches"]),
                "what": "Filter x nodes (16 threads) -> PostFilter -> Score x feasible -> NormalizeScore through the C++ plugin mirror; one device pass + three row fetches per cycle"}
    except Exception as e:  # the headline must not die with the harness
        return {"error": repr(e)[:200]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c3", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--nodes", type=int, default=0, help="override nodes per GPU")
    ap.add_argument("--pods", type=int, default=0, help="override pods per batch")
    ap.add_argument("--mode", default="full", choices=["full", "decide"])
    ap.add_argument("--force-direct", action="store_true")
    ap.add_argument("--shard", default="nodes", choices=["nodes", "pods"],
                    help="N>1: nodes = each GPU owns a slice of the cluster, pods replicated, three small exchanges (BASELINE config 4); "
                         "pods = snapshot replicated, each GPU takes its own pods, no exchange at all")
    ap.add_argument("--raw32", action="store_true", help="FULL mode with the Score matrix stored as uint32 (YODA_EVAL_RAW32, W = 5.125 B)")
    ap.add_argument("--round-values", action="store_true", help="pod variant with round scv/memory values (few requirement classes)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-c4", action="store_true", help="skip the config-4 strong-scaling block (1M nodes / N per GPU, 10k pods)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from yoda_scheduler_b200 import api, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    pod_sharded = args.shard == "pods" and world > 1
    n_nodes, n_pods, pod_seed = workload(args, rank, 1 if pod_sharded else world)   # per-GPU shard
    node_base = 0 if pod_sharded else rank * n_nodes
    snap = synth.make_snapshot(n_nodes, node_offset=node_base)
    reqs_np = synth.make_requests(n_pods, pod_seed + (rank if pod_sharded else 0), args.round_values)
    flags = (api.EVAL_FULL if args.mode == "full" else api.EVAL_DECIDE) | (api.EVAL_FORCE_DIRECT if args.force_direct else 0)
    if args.raw32 and args.mode == "full":
        flags |= api.EVAL_RAW32
    w_bytes = W_BYTES_RAW32 if (flags & api.EVAL_RAW32) else W_BYTES

    ctx = api.Context(n_nodes, device=local_rank, node_base=node_base, shard_rank=0 if pod_sharded else rank, shard_count=1 if pod_sharded else world)
    if world > 1 and not pod_sharded:
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            buf = np.zeros(128, np.uint8)
            api.check(ctx.L.yoda_comm_unique_id(buf.ctypes.data))
            idt.copy_(torch.from_numpy(buf))
        dist.broadcast(idt, 0)
        buf = idt.cpu().numpy().copy()
        api.check(ctx.L.yoda_comm_init(ctx.h, buf.ctypes.data, rank, world))
    ctx.load_snapshot(snap)

    # inputs: pinned host copy (e2e) and a device-resident copy (value)
    h_reqs = torch.empty(reqs_np.nbytes, dtype=torch.uint8, pin_memory=True)
    h_reqs.numpy()[:] = reqs_np.view(np.uint8)
    h_view = h_reqs.numpy().view(api.POD_REQ)
    d_reqs = h_reqs.to(dev)
    h_out = torch.empty(n_pods * api.POD_RES.itemsize, dtype=torch.uint8, pin_memory=True)

    stream = torch.cuda.Stream(device=dev)
    api.check(ctx.L.yoda_set_stream(ctx.h, stream.cuda_stream))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # config 5 (metric churn): every step first applies one tick -- 10% of the nodes get new per-card
    # FreeMemory / health / FreeMemorySum through the pinned update ring -- then schedules the batch
    churn = args.config == "c5"
    tick_ms = [0.0, 0]
    if churn:
        n_touch = max(1, n_nodes // 10)
        ticks = []
        card = np.arange(8, dtype=np.uint64)[None, :]
        for k in range(8):
            idx = np.sort(np.random.default_rng(k).choice(n_nodes, n_touch, replace=False)).astype(np.uint32)
            idx = idx[snap.has_scv[idx] == 1]
            gidx = (idx.astype(np.uint64) + np.uint64(node_base))[:, None]
            live = card < snap.n_cards[idx].astype(np.uint64)[:, None]
            total = snap.total_memory[idx]
            r = synth.rnd(synth.SEED + 1000 + k, synth.S_FREE, gidx, card) >> np.uint64(32)
            free = np.where(live, (total * r) >> np.uint64(32), np.uint64(0))      # new FreeMemory, <= TotalMemory
            healthy = (synth.rnd(synth.SEED + 1000 + k, synth.S_HEALTH, gidx, card) % np.uint64(64) != 0) & live
            recs = snap.records(idx)
            recs["free_memory"] = free
            recs["healthy_mask"] = (healthy.astype(np.uint64) << card).sum(axis=1).astype(np.uint32)
            recs["free_memory_sum"] = free.sum(axis=1, dtype=np.uint64)
            ticks.append((idx, recs))
        tick_no = [0]

        def tick():
            t0 = time.perf_counter()
            idx, recs = ticks[tick_no[0] % len(ticks)]
            tick_no[0] += 1
            ctx.upsert_many(idx, recs)
            ctx.commit()
            tick_ms[0] += (time.perf_counter() - t0) * 1e3
            tick_ms[1] += 1
    else:
        def tick():
            pass

    def step_device():
        tick()
        ctx.eval_device(d_reqs.data_ptr(), n_pods, flags)

    def step_e2e():
        tick()
        # node-sharded: after the exchanges every shard holds the same per-pod results; rank 0 reads them, the others pass NULL
        api.check(ctx.L.yoda_eval_batch(ctx.h, h_view.ctypes.data, n_pods, flags, h_out.data_ptr() if (rank == 0 or pod_sharded) else None))

    for _ in range(args.warmup):
        step_device()
    ctx.sync()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # ---- timed region: K steps with inputs resident in HBM ------------------------------------
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    phase = {"prep": 0.0, "max": 0.0, "score": 0.0, "write": 0.0, "total": 0.0}
    launches = 0
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(args.steps):
            step_device()
            t = ctx.timing()   # per-phase CUDA events of this step (on the same stream)
            for k in phase:
                phase[k] += getattr(t, k + "_ms")
            launches += t.kernels
        e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1) / max(args.steps, 1)
    t_last = ctx.timing()
    # ---- e2e: host buffers in, per-pod results out, through the C ABI -------------------------
    for _ in range(min(args.warmup, 2)):
        step_e2e()
    barrier()
    w0 = time.perf_counter()
    e2e_parts = [0.0, 0.0, 0.0]   # device-side H2D, kernels (first to last), D2H of each e2e step (CUDA events of the library)
    for _ in range(args.steps):
        step_e2e()
        t = ctx.timing()
        e2e_parts[0] += t.h2d_ms; e2e_parts[1] += t.total_ms; e2e_parts[2] += t.d2h_ms
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - w0) * 1e3 / max(args.steps, 1)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    e2e_parts = [v / max(args.steps, 1) for v in e2e_parts]
    if world > 1:
        tt = torch.tensor([ms, e2e_ms] + e2e_parts, device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms, e2e_ms, *e2e_parts = tt.tolist()
    total_nodes = n_nodes if pod_sharded else n_nodes * world
    total_pods = n_pods * world if pod_sharded else n_pods
    evals = float(total_pods) * total_nodes
    value = evals / (ms * 1e-3)
    e2e_value = evals / (e2e_ms * 1e-3)

    if rank == 0:
        peak, peak_src = measured_peak()
        steps = max(args.steps, 1)
        full = args.mode == "full"
        # dominant kernel = the output pass (Filter+Score+NormalizeScore -> matrices); per launch it
        # moves the whole batch's matrices for this GPU's shard
        kern_ms = phase["write"] / steps
        kern_bytes = n_pods * n_nodes * (w_bytes if full else 0.0) + n_nodes * R_NODE + n_pods * (R_POD + R_OUT)
        achieved = kern_bytes / (kern_ms * 1e-3) / 1e9 if kern_ms > 0 else 0.0
        step_bytes = algorithmic_bytes(n_pods, n_nodes, w_bytes) if full else n_nodes * R_NODE + n_pods * (R_POD + R_OUT)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if (args.config == "c4" and not pod_sharded) else "weak", "vs_baseline": None, "dtype": "u64" if t_last.path == 0 else "u32",
            "data": "synthetic",
            "config": {"workload": (f"{args.config}: {n_nodes} nodes x 8 cards replicated on every GPU, {n_pods} pods per GPU ({total_pods} total, pod-sharded, no exchange), mode={args.mode}"
                                    if pod_sharded else
                                    f"{args.config}: {n_nodes} nodes x 8 cards per GPU ({total_nodes} total, node-sharded), {n_pods} pods, mode={args.mode}")
                                   + (", round scv/memory values" if args.round_values else "")
                                   + (", Score matrix as uint32 (W = 5.125 B per evaluation)" if flags & api.EVAL_RAW32 else ""),
                       "l2": "every step writes the result matrices (>> 126 MB L2), which sweeps L2 between iterations; the snapshot is L2-resident by design",
                       "path": "narrow-fast" if t_last.path == 1 else "direct-u64", "classes": t_last.n_classes, "groups": t_last.n_groups},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(reqs_np.nbytes), "d2h_bytes_per_step": int(n_pods * api.POD_RES.itemsize),
                    "d2h_note": "per-pod results (72 B per pod) read by rank 0; node shards hold identical copies and do not read theirs back",
                    "breakdown_ms": {"h2d": e2e_parts[0], "kernels": e2e_parts[1], "d2h": e2e_parts[2],
                                     "note": "CUDA events of the library inside the e2e steps, max over ranks; the rest of ms_per_step is host time (launches, read-backs, waiting for the slowest rank)"}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "output pass (write_ms)", "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None if flags & api.EVAL_RAW32 else profiled_traffic(args.config, args.mode, t_last.path),
                         "traffic_note": "dram bytes of the output pass from the committed ncu --set full capture of this workload (profiles/traffic.json), not measured in this run",
                         "algorithmic_bytes": kern_bytes,
                         # the whole step against the same peak, from the one number the driver times (ms_per_step)
                         "step_achieved": step_bytes / (ms * 1e-3) / 1e9 if ms > 0 else 0.0,
                         "step_frac": step_bytes / (ms * 1e-3) / 1e9 / peak if ms > 0 else 0.0,
                         "frac_of_nominal_8TBs": achieved / 8000.0},
            "phases_ms": {k: v / steps for k, v in phase.items()},
        }
        if churn:
            line["churn"] = {"nodes_updated_per_tick": int(n_touch), "upsert_commit_ms_per_tick": tick_ms[0] / max(tick_ms[1], 1),
                             "bytes_per_tick": int(n_touch * 432), "note": "tick (pinned ring -> async H2D -> scatter + derive kernels -> epoch) is inside every timed step"}
        if world == 1 and args.mode == "full" and not args.no_latency:
            # the same batch in DECIDE mode (per-pod results only: no matrices, one fused score+choice pass)
            dflags = api.EVAL_DECIDE | (api.EVAL_FORCE_DIRECT if args.force_direct else 0)
            for _ in range(3):
                ctx.eval_device(d_reqs.data_ptr(), n_pods, dflags)
            ctx.sync()
            d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                d0.record(stream)
                for _ in range(10):
                    ctx.eval_device(d_reqs.data_ptr(), n_pods, dflags)
                d1.record(stream)
            torch.cuda.synchronize()
            dms = d0.elapsed_time(d1) / 10
            line["decide_mode"] = {"ms_per_step": dms, "value": evals / (dms * 1e-3), "unit": UNIT,
                                   "note": "YODA_EVAL_DECIDE: maxima, lowest/highest, feasible count and chosen node per pod; no matrices"}
        if world == 1 and args.mode == "full" and not args.no_latency and not (flags & api.EVAL_RAW32) and t_last.path == 1:
            # the same batch with the Score matrix stored as uint32 (YODA_EVAL_RAW32): W = 5.125 B per evaluation
            rflags = flags | api.EVAL_RAW32
            for _ in range(3):
                ctx.eval_device(d_reqs.data_ptr(), n_pods, rflags)
            ctx.sync()
            r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            wsum = 0.0
            with torch.cuda.stream(stream):
                r0.record(stream)
                for _ in range(10):
                    ctx.eval_device(d_reqs.data_ptr(), n_pods, rflags)
                    wsum += ctx.timing().write_ms
                r1.record(stream)
            torch.cuda.synchronize()
            rms = r0.elapsed_time(r1) / 10
            rbytes = n_pods * n_nodes * W_BYTES_RAW32 + n_nodes * R_NODE + n_pods * (R_POD + R_OUT)
            line["raw32_mode"] = {"ms_per_step": rms, "value": evals / (rms * 1e-3), "unit": UNIT, "write_ms": wsum / 10,
                                  "output_pass_GBps": rbytes / (wsum / 10 * 1e-3) / 1e9, "output_pass_frac_of_peak": rbytes / (wsum / 10 * 1e-3) / 1e9 / peak,
                                  "note": "YODA_EVAL_FULL | YODA_EVAL_RAW32: identical scores stored as uint32, W = 5.125 B per evaluation"}
            ctx.eval_device(d_reqs.data_ptr(), n_pods, flags)   # leave the int64 result behind for what follows
            ctx.sync()
        if world == 1 and not args.no_latency:
            # p50/p99 batch latency at this snapshot size, host buffers in -> per-pod results out
            lat = {}
            for pp, reps in ((1, 300), (1000, 300), (n_pods, 200)):
                if pp > n_pods:
                    continue
                sub = np.ascontiguousarray(reqs_np[:pp])
                outb = np.zeros(pp, api.POD_RES)
                ts = []
                for i in range(reps + 20):
                    t0 = time.perf_counter()
                    api.check(ctx.L.yoda_eval_batch(ctx.h, sub.ctypes.data, pp, flags, outb.ctypes.data))
                    if i >= 20:
                        ts.append((time.perf_counter() - t0) * 1e3)
                lat[f"pods={pp}"] = {"p50_ms": float(np.percentile(ts, 50)), "p99_ms": float(np.percentile(ts, 99)), "batches": reps}
            line["latency"] = {"nodes": n_nodes, "mode": args.mode, **lat}
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            all_labels = synth.make_pod_labels(min(n_pods, 4096), pod_seed, args.round_values)
            labels_fn = lambda k: [all_labels[i % len(all_labels)] for i in range(k)]
            cl, sample, _ = cpu_arm(snap, labels_fn, n_nodes, cores, 12.0)
            t0 = time.perf_counter()
            cl.eval(labels_fn(sample), faithful=True, threads=cores, full=False)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": sample * n_nodes / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{sample} pods x {n_nodes} nodes, oracle faithful driver, {cores} threads, {dt:.1f} s"}
            # faithful (per-pair label lookup + Atoi + string compare + Scv copy) and lean drivers,
            # one thread and min(16, cores) threads (16 = the framework's default parallelism); ~2 s each
            variants = {}
            for faithful in (True, False):
                for th in sorted({1, min(16, cores)}):
                    _, smp, _ = cpu_arm(snap, labels_fn, n_nodes, th, 2.0, faithful=faithful, cl=cl)
                    t0 = time.perf_counter()
                    cl.eval(labels_fn(smp), faithful=faithful, threads=th, full=False)
                    d = time.perf_counter() - t0
                    variants[f"{'faithful' if faithful else 'lean'}_{th}t"] = {"value": smp * n_nodes / d, "sample_pods": smp, "seconds": d}
            line["cpu_baseline"]["variants"] = variants
    # ---- everything below runs after the timed regions --------------------------------------------------------
    extras = {}
    if world > 1 and not pod_sharded and args.config != "c4":
        extras["parity_check"] = sharded_parity_check(ctx, api, synth, dist, torch, dev, rank, world, n_nodes, reqs_np, flags)
    if world == 1 and args.config == "c3" and args.mode == "full" and not args.no_latency:
        extras.update(fallback_paths(ctx, api, synth, torch, stream, d_reqs, snap, n_nodes, n_pods, reqs_np, flags, ms))
    ctx.close()
    del ctx
    if args.config == "c3" and args.mode == "full" and not pod_sharded and not args.no_c4:
        extras["c4_strong"] = c4_strong(args, api, synth, dist, torch, dev, rank, world, local_rank)
    if rank == 0:
        if world == 1 and not args.no_latency:
            pc = plugin_cycle(n_nodes)
            if pc:
                extras["plugin_cycle"] = pc
        line.update(extras)
        print(json.dumps(line))
    bad = extras.get("parity_check", {}).get("equal") is False
    if world > 1:
        dist.destroy_process_group()
    if bad:
        raise SystemExit("bench.py: the node-sharded result differs from the unsharded one")


if __name__ == "__main__":
    main()
