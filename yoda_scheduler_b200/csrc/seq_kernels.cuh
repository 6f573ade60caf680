// seq_kernels.cuh -- sequential-commit batch semantics.
//
// kube-scheduler handles one pod per cycle and "assumes" it onto the chosen node before the next
// cycle starts, so the next pod sees that node's NodeInfo.Pods grown by one.  In the reference the
// only arithmetic that reads NodeInfo.Pods is CalculateAllocateScore (algorithm.go:74-87): the sum
// of the running pods' scv/memory labels.  Filter, the maxima and the per-card scores never change
// with a commit, so the requirement groups, classes and class tables of the frozen-snapshot path
// stay valid; only the chosen node's static score (allocate + actual) has to be refreshed.
//
// One cooperative kernel walks the pods in queue order.  Every thread owns a fixed set of nodes
// (grid-stride) and keeps their running allocateMemorySum / static score in a private scratch
// copy, so a commit is a write by the owning thread that nobody else ever reads: one grid-wide
// barrier per pod (after the choice is reduced) is all the synchronisation there is.
#pragma once
#include <cooperative_groups.h>

#include "fast_kernels.cuh"

namespace cg = cooperative_groups;

#define SEQ_THREADS 256

// sorted position of every pod of the batch (the hot records live in sorted order)
__global__ void k_seq_positions(const PodS *__restrict__ pods, u32 P, u32 *pos_of_orig) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P) pos_of_orig[pods[i].orig] = i;
}

// private copy of the per-node state a commit changes
__global__ void k_seq_init(DevSnap s, u64 *alloc, u32 *stat) {
    u32 n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= s.n_nodes) return;
    alloc[n] = s.alloc_sum[n];
    stat[n] = s.stat32[n];
}

// CalculateAllocateScore + CalculateActualScore (algorithm.go:70-87) for one node after a commit
__device__ __forceinline__ u64 seq_static_score(u64 total, u64 fsum, u64 a) {
    const u64 t = total ? total : 1;
    const u64 actual = (fsum * 100ull / t) * 2ull;
    const u64 alloc = (total < a) ? 0ull : ((total - a) * 100ull / t * 3ull);
    return alloc + actual;
}

// status word: 0 ok, 1 = a refreshed static score left the narrow range (caller must not trust the result)
__global__ void __launch_bounds__(SEQ_THREADS) k_seq_commit(DevSnap s, const PodS *__restrict__ pods, const PodC *__restrict__ podc,
                                                            const u32 *__restrict__ pos_of_orig, const u32 *__restrict__ order, u32 P,
                                                            const u32 *__restrict__ tables, u64 *alloc, u32 *stat, u64 *hilo, u64 *count,
                                                            u64 *dkey, u32 *status) {
    cg::grid_group grid = cg::this_grid();
    __shared__ u32 s_hi[SEQ_THREADS / 32], s_lo[SEQ_THREADS / 32], s_cnt[SEQ_THREADS / 32], s_best[SEQ_THREADS / 32];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u32 tid = blockIdx.x * SEQ_THREADS + threadIdx.x, nthreads = gridDim.x * SEQ_THREADS;
    for (u32 k = 0; k < P; ++k) {
        const u32 orig = order ? order[k] : k;
        const uint4 a = reinterpret_cast<const uint4 *>(pods)[2 * pos_of_orig[orig]];
        const uint4 b = reinterpret_cast<const uint4 *>(pods)[2 * pos_of_orig[orig] + 1];
        const u32 m = a.x, c = a.y, num = a.z, reqM = a.w & 0xFFu, reqC = (a.w >> 8) & 0xFFu, cls = b.x;
        const bool has_clock = (a.w >> 16) & YODA_POD_HAS_CLOCK;
        // this thread's nodes: best feasible score (lowest index on ties), lowest score, feasible count
        u32 hi = 0, lo = 0xFFFFFFFFu, cnt = 0, best = 0xFFFFFFFFu;
        for (u32 n0 = blockIdx.x * SEQ_THREADS; n0 < s.n_nodes; n0 += nthreads) {  // warp-uniform trip count
            const u32 n = n0 + threadIdx.x;
            const bool in = n < s.n_nodes;
            const u32 nn = in ? n : 0;
            bool fit;
            u32 bound;
            const u32 raw = hot_eval_mem(s, tables, nn, cls, m, num, reqM, has_clock, c, reqC, in && (s.meta[nn] & META_HAS_SCV),
                                         s.cardnum32[nn], in ? stat[nn] : 0u, fit, bound);
            if (fit) {
                cnt++;
                lo = min(lo, raw);
                if (best == 0xFFFFFFFFu || raw > hi) { hi = raw; best = n; }  // n ascends per thread: ties keep the lower index
            }
        }
        // warp, then CTA, then grid: highest with the lowest index, lowest, count
        const u32 whi = __reduce_max_sync(0xFFFFFFFFu, best == 0xFFFFFFFFu ? 0u : hi);
        const u32 wbest = __reduce_min_sync(0xFFFFFFFFu, (best != 0xFFFFFFFFu && hi == whi) ? best : 0xFFFFFFFFu);
        const u32 wlo = __reduce_min_sync(0xFFFFFFFFu, lo);
        const u32 wcnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
        if (lane == 0) { s_hi[warp] = whi; s_lo[warp] = wlo; s_cnt[warp] = wcnt; s_best[warp] = wbest; }
        __syncthreads();
        if (threadIdx.x == 0) {
            u32 bhi = 0, blo = 0xFFFFFFFFu, bcnt = 0, bbest = 0xFFFFFFFFu;
#pragma unroll
            for (int w = 0; w < SEQ_THREADS / 32; ++w) {
                if (s_cnt[w]) {
                    if (bbest == 0xFFFFFFFFu || s_hi[w] > bhi || (s_hi[w] == bhi && s_best[w] < bbest)) { bhi = s_hi[w]; bbest = s_best[w]; }
                    blo = min(blo, s_lo[w]);
                    bcnt += s_cnt[w];
                }
            }
            if (bcnt) {
                atomicMax(&hilo[2 * (size_t)orig], (u64)bhi);
                atomicMax(&hilo[2 * (size_t)orig + 1], ~(u64)blo);
                atomicAdd(&count[orig], (u64)bcnt);
                atomicMax(&dkey[orig], ((u64)bhi << 32) | (u64)(0xFFFFFFFFu - (s.node_base + bbest)));
            }
        }
        grid.sync();
        // commit: the owner of the chosen node grows its allocateMemorySum (algorithm.go:75-80) and refreshes
        // the static score; nobody else reads these two entries
        const PodC pc = podc[orig];
        if ((pc.flags & YODA_POD_HAS_MEMORY) && __ldcg(&count[orig]) != 0) {
            const u32 chosen = 0xFFFFFFFFu - (u32)__ldcg(&dkey[orig]) - s.node_base;
            if (chosen < s.n_nodes && (chosen % nthreads) == tid) {
                const u64 na = alloc[chosen] + pc.mem;  // wraps like Go's uint64
                alloc[chosen] = na;
                const u64 st = seq_static_score(s.total_sum[chosen], s.free_sum[chosen], na);
                if (st >= (1ull << 24)) atomicOr(status, 1u);
                stat[chosen] = (u32)st;
            }
        }
    }
}
