// tail_kernels.cuh -- what kube-scheduler v1.20 does AROUND the plugin for each pod, on the device, over the
// matrices a YODA_EVAL_FULL batch left in HBM (oracle: ora_framework_tail):
//
//   numFeasibleNodesToFind   how many feasible nodes a cycle looks for (percentageOfNodesToScore, 0 = adaptive,
//                            deploy/yoda-scheduler.yaml:18)                                         -- host, runtime.cu
//   findNodesThatPassFilters nodes visited round-robin from nextStartNodeIndex until that many passed Filter;
//                            the next cycle starts where this one stopped                          -- k_tail_window
//   prioritizeNodes          Yoda.NormalizeScore (scheduler.go:132-157) over the nodes FOUND, times the plugin weight
//                            (300, deploy/yoda-scheduler.yaml:30), plus the other plugins' scores  -- k_tail_pick
//   selectHost               highest total, ties by reservoir sampling with a seeded generator      -- k_tail_pick
//
// Upstream visits the nodes from 16 goroutines, so which nodes make the cut is a race there; here (and in the oracle)
// they are taken strictly in visiting order -- one of the outcomes that race can produce, and a reproducible one.
// The feasible bits and the Score rows are exactly the ones the batch evaluation wrote; nothing is re-scored.
#pragma once
#include "common.cuh"
#include "direct_kernels.cuh"

struct TailOut {  // == yoda_tail_result
    i32 chosen;
    u32 sampled, processed, start;
    i64 lowest, highest, top;
    u32 ties, pad;
};

#define TAIL_TILE 1024  // nodes per popcount tile (32 words of the feasible-bit row)

// feasible nodes per tile of every pod's bit row
__global__ void k_tail_tilecnt(const u32 *__restrict__ bits, u64 words_per_row, u32 P, u32 tiles, u32 *tilecnt) {
    const u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (u64)P * tiles) return;
    const u32 p = (u32)(t / tiles), tile = (u32)(t % tiles);
    const uint4 *w = reinterpret_cast<const uint4 *>(bits + (size_t)p * words_per_row + (size_t)tile * 32);
    u32 c = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint4 v = w[i];
        c += __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
    }
    tilecnt[t] = c;
}

__device__ __forceinline__ u64 tail_sm64(u64 x) {
    u64 z = x + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ u64 tail_rand(u64 seed, u64 pod, u64 k) { return tail_sm64(tail_sm64(tail_sm64(seed) + pod) + k); }

// warp-wide helpers over one pod's bit row --------------------------------------------------------------------------
// feasible nodes with index < n
__device__ __forceinline__ u32 tail_rank(const u32 *__restrict__ row, const u32 *__restrict__ tc, u32 n, u32 lane) {
    const u32 tile = n / TAIL_TILE;
    u32 c = 0;
    for (u32 t = lane; t < tile; t += 32) c += tc[t];
    const u32 w = tile * 32 + lane, first = w * 32;  // lane's word of the tile holding n
    if (first < n) {
        const u32 v = row[w];
        c += (n - first >= 32) ? __popc(v) : __popc(v & ((1u << (n - first)) - 1u));
    }
    return __reduce_add_sync(0xFFFFFFFFu, c);
}
// index of the r-th feasible node (r >= 1, r <= feasible nodes of the row)
__device__ __forceinline__ u32 tail_select(const u32 *__restrict__ row, const u32 *__restrict__ tc, u32 tiles, u32 r, u32 lane) {
    u32 before = 0, tile = 0;
    for (u32 t0 = 0; t0 < tiles; t0 += 32) {  // the tile in which the running count reaches r
        const u32 c = (t0 + lane < tiles) ? tc[t0 + lane] : 0u;
        u32 inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 v = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if ((int)lane >= o) inc += v;
        }
        const u32 hit = __ballot_sync(0xFFFFFFFFu, before + inc >= r);
        if (hit) {
            const u32 l = __ffs(hit) - 1;
            tile = t0 + l;
            before += __shfl_sync(0xFFFFFFFFu, inc - c, l);
            break;
        }
        before += __shfl_sync(0xFFFFFFFFu, inc, 31);
    }
    const u32 v = row[tile * 32 + lane];  // then the word, then the bit
    const u32 c = __popc(v);
    u32 inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u32 x = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if ((int)lane >= o) inc += x;
    }
    const u32 hit = __ballot_sync(0xFFFFFFFFu, before + inc >= r);
    const u32 l = __ffs(hit) - 1;
    const u32 word = __shfl_sync(0xFFFFFFFFu, v, l);
    u32 need = r - (before + __shfl_sync(0xFFFFFFFFu, inc - c, l));  // 1-based inside the word
    u32 wv = word;
    for (u32 i = 1; i < need; ++i) wv &= wv - 1;  // drop the lowest set bits before it
    return (tile * 32 + l) * 32 + (__ffs(wv) - 1);
}

// findNodesThatPassFilters for every pod.  chain: ONE warp walks the pods in order, each starting where the previous
// one stopped (that is what successive scheduling cycles do); otherwise one warp per pod, all from start_index.
__global__ void k_tail_window(const u32 *__restrict__ bits, u64 words_per_row, const u32 *__restrict__ tilecnt, u32 tiles, u32 P, u32 N, u32 want,
                              u32 start_index, int chain, TailOut *out) {
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    u32 p0 = chain ? 0 : warp, p1 = chain ? P : min(P, warp + 1);
    if (chain && warp != 0) return;
    u32 next = N ? start_index % N : 0;
    for (u32 p = p0; p < p1; ++p) {
        const u32 *row = bits + (size_t)p * words_per_row;
        const u32 *tc = tilecnt + (size_t)p * tiles;
        const u32 s0 = chain ? next : (N ? start_index % N : 0);
        u32 total = 0;
        for (u32 t = lane; t < tiles; t += 32) total += tc[t];
        total = __reduce_add_sync(0xFFFFFFFFu, total);
        u32 visited = N, nf = total;
        if (total >= want && want > 0) {
            const u32 before = tail_rank(row, tc, s0, lane);
            const bool wraps = total - before < want;
            const u32 idx = tail_select(row, tc, tiles, wraps ? want - (total - before) : before + want, lane);
            visited = wraps ? (N - s0) + idx + 1 : idx - s0 + 1;
            nf = want;
        }
        next = N ? (s0 + visited) % N : 0;
        if (lane == 0) {
            TailOut o;
            o.chosen = -1; o.sampled = nf; o.processed = visited; o.start = s0;
            o.lowest = 0; o.highest = 0; o.top = 0; o.ties = 0; o.pad = 0;
            out[p] = o;
        }
    }
}

// prioritizeNodes + selectHost for one pod per warp, over the window k_tail_window found.
template <bool RAW32>
__global__ void k_tail_pick(const u32 *__restrict__ bits, u64 words_per_row, const void *__restrict__ raw_rows, u64 stride, u32 P, u32 N, u32 weight,
                            u64 seed, const i64 *__restrict__ other, u32 node_base, TailOut *out) {
    const u32 lane = threadIdx.x & 31;
    const u32 p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (p >= P) return;
    TailOut o = out[p];
    if (o.sampled == 0) return;
    const u32 *row = bits + (size_t)p * words_per_row;
    const u32 s0 = o.start, V = o.processed;
    auto raw_at = [&](u32 n) -> i64 {
        return RAW32 ? (i64) reinterpret_cast<const u32 *>(raw_rows)[(size_t)p * stride + n] : reinterpret_cast<const i64 *>(raw_rows)[(size_t)p * stride + n];
    };
    auto fits = [&](u32 n) -> bool { return (row[n >> 5] >> (n & 31)) & 1u; };
    if (o.sampled == 1) {  // a single feasible node is taken without scoring
        u32 found = 0xFFFFFFFFu;
        for (u32 i = lane; i < V; i += 32) {
            u32 n = s0 + i;
            n = n >= N ? n - N : n;
            if (fits(n)) found = n;
        }
        found = __reduce_min_sync(0xFFFFFFFFu, found);
        if (lane == 0) {
            o.chosen = (i32)(node_base + found);
            o.ties = 1;
            out[p] = o;
        }
        return;
    }
    // NormalizeScore over the nodes found: highest starts at 0, lowest at the first score (scheduler.go:134-149)
    i64 hi = 0, lo = INT64_MAX;
    for (u32 i = lane; i < V; i += 32) {
        u32 n = s0 + i;
        n = n >= N ? n - N : n;
        if (fits(n)) {
            const i64 r = raw_at(n);
            hi = max(hi, r);
            lo = min(lo, r);
        }
    }
    for (int k = 16; k; k >>= 1) {
        hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, k));
        lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, k));
    }
    if (hi == lo) lo = (i64)((u64)lo - 1ull);
    auto total_of = [&](u32 n) -> i64 {
        const i64 nv = dev_norm_i64(raw_at(n), lo, hi);  // scheduler.go:152
        return (i64)((u64)nv * (u64)weight + (u64)(other ? other[n] : 0));
    };
    i64 best = INT64_MIN;
    for (u32 i = lane; i < V; i += 32) {
        u32 n = s0 + i;
        n = n >= N ? n - N : n;
        if (fits(n)) best = max(best, total_of(n));
    }
    for (int k = 16; k; k >>= 1) best = max(best, __shfl_xor_sync(0xFFFFFFFFu, best, k));
    // selectHost: in visiting order, the t-th node holding the top total replaces the candidate iff rand(t) % t == 0
    u32 cnt = 0, chosen = 0xFFFFFFFFu;
    for (u32 i0 = 0; i0 < V; i0 += 32) {
        const u32 i = i0 + lane;
        u32 n = s0 + i;
        n = n >= N ? n - N : n;
        const bool tie = i < V && fits(n) && total_of(n) == best;
        const u32 m = __ballot_sync(0xFFFFFFFFu, tie);
        if (!m) continue;
        const u32 t = cnt + __popc(m & ((1u << lane) - 1u)) + 1;  // this lane's ordinal among the ties, if it is one
        const bool takes = tie && (t == 1 || tail_rand(seed, p, t) % t == 0);
        const u32 tm = __ballot_sync(0xFFFFFFFFu, takes);
        if (tm) chosen = __shfl_sync(0xFFFFFFFFu, n, 31 - __clz(tm));  // the last taker of this stretch
        cnt += __popc(m);
    }
    if (lane == 0) {
        o.chosen = (i32)(node_base + chosen);
        o.lowest = lo; o.highest = hi; o.top = best; o.ties = cnt;
        out[p] = o;
    }
}
