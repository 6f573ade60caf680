// direct_kernels.cuh -- exact-u64 "direct" evaluation kernels and the snapshot maintenance
// kernels.  The direct kernels evaluate every (pod, node) pair literally, with real 64-bit
// divisions and Go's wrap-around, for ANY uint64 input.  They are the path for snapshots or
// pods outside the narrow (uint32) range and the on-device cross-check of the fast path.
//
// Reference arithmetic restated here (paths relative to the reference checkout):
//   filter.go:11-58 (predicates), collection.go:30-78 (maxima), algorithm.go:28-87 (score),
//   filter.go:84-86 (uint64 -> int64), scheduler.go:132-157 (NormalizeScore).
#pragma once
#include "common.cuh"

// ------------------------------------------------------------------------------------------
// snapshot maintenance
// ------------------------------------------------------------------------------------------

// Scatter staged update records into the wide planes.  One thread per (record, card slot).
__global__ void k_scatter_updates(DevSnap s, const UpdRec *__restrict__ upd, u32 n_upd) {
    u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    u32 r = t / G, g = t % G;
    if (r >= n_upd) return;
    const UpdRec &u = upd[r];
    const u32 n = u.idx;
    if (u.op & UPD_REC) {
        const bool live = g < u.rec.n_cards;
        s.f[F_FREE][(size_t)n * G + g] = live ? u.rec.free_memory[g] : 0;
        s.f[F_TOTAL][(size_t)n * G + g] = live ? u.rec.total_memory[g] : 0;
        s.f[F_CLOCK][(size_t)n * G + g] = live ? u.rec.clock[g] : 0;
        s.f[F_CORE][(size_t)n * G + g] = live ? u.rec.core[g] : 0;
        s.f[F_BW][(size_t)n * G + g] = live ? u.rec.bandwidth[g] : 0;
        s.f[F_POWER][(size_t)n * G + g] = live ? u.rec.power[g] : 0;
        if (g == 0) {
            s.card_number[n] = u.rec.card_number;
            s.free_sum[n] = u.rec.free_memory_sum;
            s.total_sum[n] = u.rec.total_memory_sum;
            const u32 nc = u.rec.n_cards;
            const u32 hm = u.rec.healthy_mask & ((1u << nc) - 1u) & 0xFFu;
            s.meta[n] = META_HAS_SCV | (nc << 8) | hm | (s.meta[n] & META_WIDE);  // the derive step below settles the wide bit
        }
    }
    if (g == 0) {
        if (u.op & UPD_ALLOC) s.alloc_sum[n] = u.alloc;
        if (u.op & UPD_REMOVE) s.meta[n] &= META_WIDE;
    }
}

// Derived per-node data, recomputed for every node an update record touched:
//   stat[n]  = CalculateAllocateScore + CalculateActualScore  (algorithm.go:70-87), wrapping u64
//   narrow planes + the snapshot-wide "narrow eligible" verdict:
//     every card field < 2^32-1, CardNumber < 2^32-1, stat < 2^24.
// wide_flag[0] counts the nodes with an SCV that fall outside the narrow range (META_WIDE marks them),
// wide_flag[1] if a node's ordering key (see DevSnap::order) changed.
#define NARROW_FIELD_LIMIT 0xFFFFFFFFull  /* exclusive */
#define NARROW_STAT_LIMIT (1ull << 24)    /* exclusive */
__device__ __forceinline__ void prepare_node(const DevSnap &s, u32 n, u32 *wide_flag) {
    const u32 meta = s.meta[n];
    bool wide = false;
    u64 st = 0;
    if (meta & META_HAS_SCV) {
        const u64 total = s.total_sum[n], fsum = s.free_sum[n], a = s.alloc_sum[n];
        // total > 0 is enforced at upsert (YODA_E_DIV0); guard anyway so the device never traps
        const u64 t = total ? total : 1;
        const u64 actual = (fsum * 100ull / t) * 2ull;                          // algorithm.go:71
        const u64 alloc = (total < a) ? 0ull : ((total - a) * 100ull / t * 3ull);  // algorithm.go:82-86
        st = alloc + actual;
        wide |= st >= NARROW_STAT_LIMIT;
        wide |= s.card_number[n] >= NARROW_FIELD_LIMIT;
    }
    s.stat[n] = st;
    s.stat32[n] = (u32)st;
    s.cardnum32[n] = (u32)s.card_number[n];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const size_t o = (size_t)n * G + g;
        u64 any = 0;
#pragma unroll
        for (int f = 0; f < NFIELD; ++f) any |= (s.f[f][o] >= NARROW_FIELD_LIMIT) ? 1ull : 0ull;
        wide |= (any != 0) && (meta & META_HAS_SCV);
#pragma unroll
        for (int f = 0; f < NFIELD; ++f) s.n32[f][o] = (u32)s.f[f][o];
    }
    // snapshot-wide verdict = number of wide nodes; this node's contribution changes by (now wide) - (was wide)
    const bool was = (meta & META_WIDE) != 0;
    if (wide != was) atomicAdd(wide_flag, wide ? 1u : 0xFFFFFFFFu);
    // cards worth visiting for the maxima, and the k-th largest FreeMemory of the healthy cards
    {
        const u32 nc = meta_ncards(meta), hm = meta_healthy(meta);
        u64 v[NFIELD][G];
#pragma unroll
        for (int f = 0; f < NFIELD; ++f)
#pragma unroll
            for (int g = 0; g < G; ++g) v[f][g] = s.f[f][(size_t)n * G + g];
        u32 keep = 0, tk[G];
#pragma unroll
        for (int k = 0; k < G; ++k) tk[k] = 0;
#pragma unroll
        for (int b = 0; b < G; ++b) {
            bool dominated = b >= (int)nc;
            u32 rank = 0;
#pragma unroll
            for (int a = 0; a < G; ++a) {
                if (a == b) continue;
                bool ge = a < (int)nc, gt = false;
#pragma unroll
                for (int f = 0; f < NFIELD; ++f) {
                    ge = ge && v[f][a] >= v[f][b];
                    gt = gt || v[f][a] > v[f][b];
                }
                dominated = dominated || (ge && (gt || a < b));  // of identical cards the first stays
                if (((hm >> a) & 1u) && (v[F_FREE][a] > v[F_FREE][b] || (v[F_FREE][a] == v[F_FREE][b] && a < b))) rank++;
            }
            if (!dominated) keep |= 1u << b;
            if ((hm >> b) & 1u) {
#pragma unroll
                for (int k = 0; k < G; ++k)
                    if (rank == (u32)k) tk[k] = (u32)v[F_FREE][b];
            }
        }
        s.meta[n] = (meta & 0x00FFFFFFu & ~META_WIDE) | (wide ? META_WIDE : 0u) | (keep << 24);
        uint4 *dst = reinterpret_cast<uint4 *>(s.tk32 + (size_t)n * G);
        dst[0] = make_uint4(tk[0], tk[1], tk[2], tk[3]);
        dst[1] = make_uint4(tk[4], tk[5], tk[6], tk[7]);
    }
    const u32 key = (meta & META_HAS_SCV) ? min((u32)s.f[F_CLOCK][(size_t)n * G], 0xFFFFFFFDu) : 0xFFFFFFFEu;
    if (s.order_key[n] != key) {
        s.order_key[n] = key;
        wide_flag[1] = 1u;
    }
}

// one thread per staged update record, right after k_scatter_updates: a commit costs what its updates cost
__global__ void k_prepare_updates(DevSnap s, const UpdRec *__restrict__ upd, u32 n_upd, u32 *wide_flag) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_upd) prepare_node(s, upd[i].idx, wide_flag);
}

__global__ void k_iota(u32 *v, u32 n) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = i;
}

// ------------------------------------------------------------------------------------------
// per-pod records
// ------------------------------------------------------------------------------------------
__global__ void k_pod_prep(const yoda_pod_req *__restrict__ req, PodC *__restrict__ pods, u32 P) {
    u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const yoda_pod_req r = req[p];
    PodC c;
    c.flags = r.flags & 7u;
    c.num = (r.flags & YODA_POD_HAS_NUMBER) ? r.number : 1ull;
    c.mem = (r.flags & YODA_POD_HAS_MEMORY) ? r.memory : 0ull;
    c.clk = (r.flags & YODA_POD_HAS_CLOCK) ? r.clock : 0ull;
    c.pad = 0;
    pods[p] = c;
}

// identities of the partial buffers (see common.cuh)
__global__ void k_init_partials(u64 *max6, u64 *count, u64 *hilo, u64 *ckey, u32 P) {
    u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
#pragma unroll
    for (int i = 0; i < 6; ++i) max6[(size_t)p * 6 + i] = 1ull;  // collection.go:31-38
    count[p] = 0;
    hilo[2 * (size_t)p] = 0;
    hilo[2 * (size_t)p + 1] = 0;
    ckey[p] = 0;
}

// ------------------------------------------------------------------------------------------
// the per-pair predicates, shared by all direct kernels
// ------------------------------------------------------------------------------------------
// Returns Filter's verdict for (pod, node) and the bit mask of cards with
// FreeMemory >= memory && Clock >= clock (collection.go:46, algorithm.go:48 -- no health test).
__device__ __forceinline__ bool direct_fit(const PodC &pod, u32 meta, u64 card_number, const u64 *fr, const u64 *ck,
                                           u32 &qual) {
    qual = 0;
    if (!(meta & META_HAS_SCV)) return false;            // scheduler.go:81-84
    const u32 nc = meta_ncards(meta), hm = meta_healthy(meta);
    u32 cnt_m = 0, cnt_c = 0;
#pragma unroll
    for (int g = 0; g < G; ++g) {
        if (g < (int)nc) {
            const u32 h = (hm >> g) & 1u;
            const bool qm = fr[g] >= pod.mem, qc = ck[g] >= pod.clk;
            cnt_m += h & (u32)qm;                        // filter.go:52-54
            cnt_c += h & (u32)(ck[g] == pod.clk);        // filter.go:56-58
            qual |= (u32)(qm && qc) << g;
        }
    }
    bool fit = pod.num <= card_number;                    // filter.go:13,15
    if (pod.flags & YODA_POD_HAS_MEMORY) fit = fit && ((u64)cnt_m >= pod.num);  // filter.go:27
    if (pod.flags & YODA_POD_HAS_CLOCK) fit = fit && ((u64)cnt_c >= pod.num);   // filter.go:44
    return fit;
}

// algorithm.go:57-68, wrapping u64, clock divided by MaxBandwidth (:60)
__device__ __forceinline__ u64 direct_card_score(const u64 *mx, u64 bw, u64 clock, u64 core, u64 power, u64 fre,
                                                 u64 total) {
    const u64 b = bw * 100ull / mx[0];
    const u64 c = clock * 100ull / mx[0];
    const u64 co = core * 100ull / mx[2];
    const u64 pw = power * 100ull / mx[4];
    const u64 fm = fre * 100ull / mx[3];
    const u64 tm = total * 100ull / mx[5];
    return (b + c + co + pw) + fm * 2ull + tm;
}

// filter.go:84-86
__device__ __forceinline__ i64 dev_u64_to_i64(u64 v) { return (v >> 63) ? 0ll : (i64)v; }

// scheduler.go:147-153 for one score, given lowest/highest AFTER the lowest-- adjustment
__device__ __forceinline__ i64 dev_norm_i64(i64 s, i64 lo, i64 hi) {
    const i64 num = (i64)(((u64)s - (u64)lo) * 100ull);
    const i64 den = (i64)((u64)hi - (u64)lo);
    if (num == INT64_MIN && den == -1) return num;  // Go: MinInt64 / -1 == MinInt64
    return num / den;                                // truncates toward zero like Go
}
// normalized value as stored in the uint8 matrix: 0..100, or 255 when the int64 rescale wrapped
// outside [0,100] (the k8s framework rejects such a cycle)
__device__ __forceinline__ u8 dev_norm_u8(i64 v) { return (v >= 0 && v <= 100) ? (u8)v : (u8)255; }

// ------------------------------------------------------------------------------------------
// direct phase A/B: thread = pod, node tiles staged through shared memory
// ------------------------------------------------------------------------------------------
#define DTN 32   // nodes per shared-memory tile
#define DPT 128  // pods (threads) per CTA
struct DirectTile {
    u64 f[NFIELD][DTN * G];
    u64 cardnum[DTN];
    u64 stat[DTN];
    u32 meta[DTN];
    u32 node[DTN];  // local node index of every tile slot
};

// Stage nodes [n0, n1) -- positions in `node_list` when one is given (the snapshot's wide nodes, see below),
// node indices otherwise -- into shared memory.
__device__ __forceinline__ void direct_load_tile(DirectTile &sm, const DevSnap &s, const u32 *__restrict__ node_list, u32 n0, u32 n1) {
    const u32 cnt = (n1 - n0) * G;
    if (node_list) {
#pragma unroll
        for (int f = 0; f < NFIELD; ++f)
            for (u32 i = threadIdx.x; i < cnt; i += blockDim.x) sm.f[f][i] = s.f[f][(size_t)node_list[n0 + i / G] * G + i % G];
    } else {
#pragma unroll
        for (int f = 0; f < NFIELD; ++f)
            for (u32 i = threadIdx.x; i < cnt; i += blockDim.x) sm.f[f][i] = s.f[f][(size_t)n0 * G + i];
    }
    for (u32 i = threadIdx.x; i < n1 - n0; i += blockDim.x) {
        const u32 n = node_list ? node_list[n0 + i] : n0 + i;
        sm.cardnum[i] = s.card_number[n];
        sm.stat[i] = s.stat[n];
        sm.meta[i] = s.meta[n];
        sm.node[i] = n;
    }
}

// lowest/highest as NormalizeScore uses them (scheduler.go:134-149), from the reduced partials
__device__ __forceinline__ void dev_lo_hi(const u64 *hilo, u32 p, i64 &lo, i64 &hi) {
    if (hilo[2 * (size_t)p + 1] == 0) {  // identity: no feasible node on any shard
        lo = 0;
        hi = 0;
        return;
    }
    hi = (i64)hilo[2 * (size_t)p];
    lo = (i64)(~hilo[2 * (size_t)p + 1]);
    if (hi == lo) lo = (i64)((u64)lo - 1ull);
}

// Output side of the thread = pod kernel (MODE 2): where the verdicts and scores of the listed nodes go.
// Device-side path selection (node-sharded FULL batches): whether a shard's class tables fit the packed format is known
// on the device only (FastCtl::err); instead of a host round trip both paths are launched and each kernel looks at the
// word: GATE_IF_ZERO kernels (table path, wide-node list) run when it is 0, GATE_IF_SET kernels (exact path) otherwise.
#define GATE_NONE 0
#define GATE_IF_ZERO 1
#define GATE_IF_SET 2
__device__ __forceinline__ bool gate_closed(const u32 *gate, int mode) {
    if (mode == GATE_NONE) return false;
    const bool set = *reinterpret_cast<const volatile u32 *>(gate) != 0;
    return mode == GATE_IF_ZERO ? set : !set;
}

struct DirectOut {
    const u32 *gate;  // see gate_closed
    int gate_mode;
    u64 *count, *ckey;
    i64 *raw;    // FULL mode: result matrices (null otherwise); entries of the listed nodes are overwritten
    u8 *norm;
    u32 *bits;
    u64 stride;
    u32 packed;  // choice keys are (0xFFFFFFFF << 32 | ~global index): merged DECIDE after k_wide_promote
};

// MODE 0: collection.CollectMaxValues.  MODE 1: Score reduce (lowest/highest).
// MODE 2: output pass (Score, NormalizeScore, feasible bit, count, node choice) for a LIST of nodes -- the wide
// nodes of a mixed snapshot, a handful among many narrow ones, so thread = pod is the parallel axis there too.
// node_list == nullptr: every node of the snapshot; otherwise only the n_list nodes it names.
template <int MODE>
__global__ void __launch_bounds__(DPT) k_direct_pods(DevSnap s, const PodC *__restrict__ pods, u32 P, u32 nodes_per_cta,
                                                     u64 *max6, u64 *hilo, const u32 *__restrict__ node_list, u32 n_list, DirectOut out) {
    __shared__ DirectTile sm;
    if (gate_closed(out.gate, out.gate_mode)) return;
    const u32 p = blockIdx.x * DPT + threadIdx.x;
    const bool active = p < P;
    PodC pod;
    if (active) pod = pods[p];
    u64 mx[6] = {1, 1, 1, 1, 1, 1};
    u64 hi = 0, nlo = 0;
    bool any = false;
    i64 lo2 = 0, hi2 = 0;
    u32 cnt = 0, best = 0xFFFFFFFFu;
    if (MODE >= 1 && active) {
#pragma unroll
        for (int i = 0; i < 6; ++i) mx[i] = max6[(size_t)p * 6 + i];
    }
    if (MODE == 2 && active) dev_lo_hi(hilo, p, lo2, hi2);
    const u32 n_total = node_list ? n_list : s.n_nodes;
    const u32 nbeg = blockIdx.y * nodes_per_cta;
    const u32 nend = min(n_total, nbeg + nodes_per_cta);
    for (u32 n0 = nbeg; n0 < nend; n0 += DTN) {
        const u32 n1 = min(nend, n0 + DTN);
        __syncthreads();
        direct_load_tile(sm, s, node_list, n0, n1);
        __syncthreads();
        if (!active) continue;
        for (u32 i = 0; i < n1 - n0; ++i) {
            u32 qual;
            const bool fit = direct_fit(pod, sm.meta[i], sm.cardnum[i], &sm.f[F_FREE][i * G], &sm.f[F_CLOCK][i * G], qual);
            if (MODE == 2) {
                u64 basic = 0;
                if (fit)
                    for (int g = 0; g < G; ++g)
                        if ((qual >> g) & 1u)
                            basic += direct_card_score(mx, sm.f[F_BW][i * G + g], sm.f[F_CLOCK][i * G + g], sm.f[F_CORE][i * G + g],
                                                       sm.f[F_POWER][i * G + g], sm.f[F_FREE][i * G + g], sm.f[F_TOTAL][i * G + g]);
                const i64 raw = (sm.meta[i] & META_HAS_SCV) ? dev_u64_to_i64(basic + sm.stat[i]) : 0;
                const u32 n = sm.node[i];
                if (fit) {
                    cnt++;
                    if (raw == hi2) best = min(best, n);
                }
                if (out.raw) {
                    out.raw[(size_t)p * out.stride + n] = raw;
                    out.norm[(size_t)p * out.stride + n] = fit ? dev_norm_u8(dev_norm_i64(raw, lo2, hi2)) : (u8)0;
                    u32 *w = &out.bits[(size_t)p * (out.stride / 32) + (n >> 5)];
                    if (fit) atomicOr(w, 1u << (n & 31)); else atomicAnd(w, ~(1u << (n & 31)));
                }
                continue;
            }
            if (!fit) continue;
            if (MODE == 0) {
                any = true;
                for (int g = 0; g < G; ++g) {
                    if ((qual >> g) & 1u) {  // collection.go:59-78
                        mx[0] = max(mx[0], sm.f[F_BW][i * G + g]);
                        mx[1] = max(mx[1], sm.f[F_CLOCK][i * G + g]);
                        mx[2] = max(mx[2], sm.f[F_CORE][i * G + g]);
                        mx[3] = max(mx[3], sm.f[F_FREE][i * G + g]);
                        mx[4] = max(mx[4], sm.f[F_POWER][i * G + g]);
                        mx[5] = max(mx[5], sm.f[F_TOTAL][i * G + g]);
                    }
                }
            } else {
                u64 basic = 0;
                for (int g = 0; g < G; ++g)
                    if ((qual >> g) & 1u)
                        basic += direct_card_score(mx, sm.f[F_BW][i * G + g], sm.f[F_CLOCK][i * G + g], sm.f[F_CORE][i * G + g],
                                                   sm.f[F_POWER][i * G + g], sm.f[F_FREE][i * G + g], sm.f[F_TOTAL][i * G + g]);
                const i64 raw = dev_u64_to_i64(basic + sm.stat[i]);  // algorithm.go:38, scheduler.go:128
                hi = max(hi, (u64)raw);
                nlo = max(nlo, ~(u64)raw);
            }
        }
    }
    if (!active) return;
    if (MODE == 0) {
        if (any) {
#pragma unroll
            for (int i = 0; i < 6; ++i)
                if (mx[i] > 1) atomicMax(&max6[(size_t)p * 6 + i], mx[i]);
        }
    } else if (MODE == 1) {
        if (nlo) {  // at least one feasible node seen
            atomicMax(&hilo[2 * (size_t)p], hi);
            atomicMax(&hilo[2 * (size_t)p + 1], nlo);
        }
    } else {
        if (cnt) atomicAdd(&out.count[p], (u64)cnt);
        if (best != 0xFFFFFFFFu)
            atomicMax(&out.ckey[p], out.packed ? (0xFFFFFFFF00000000ull | (u64)(0xFFFFFFFFu - (s.node_base + best))) : ckey_of(s.node_base + best));
    }
}

// Merged DECIDE on a mixed snapshot: the score pass left, per pod, (best narrow score << 32 | ~index) in dkey.  Once
// `highest` is final (the wide nodes' scores included), a narrow candidate stands only if its score is that
// highest; it is re-keyed to (0xFFFFFFFF << 32 | ~index) so that the wide nodes' candidates (k_direct_pods<2>,
// packed) compete with it on the index alone.
__global__ void k_wide_promote(const u64 *__restrict__ hilo, const u64 *__restrict__ count, u64 *dkey, u32 P) {
    const u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const u64 k = dkey[p];
    const bool stands = count[p] != 0 && (k >> 32) == hilo[2 * (size_t)p];
    dkey[p] = stands ? (0xFFFFFFFF00000000ull | (k & 0xFFFFFFFFull)) : 0ull;
}

// the snapshot's wide nodes (META_WIDE), gathered at commit; order is irrelevant to every consumer
__global__ void k_wide_list(const u32 *__restrict__ meta, u32 n_nodes, u32 *list, u32 *counter, u32 cap) {
    const u32 n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < n_nodes && (meta[n] & META_WIDE)) {
        const u32 k = atomicAdd(counter, 1u);
        if (k < cap) list[k] = n;
    }
}

// ------------------------------------------------------------------------------------------
// direct phase C: thread = node (coalesced rows), a chunk of pods per CTA
// Writes (WRITE) Score for every node, the normalized score and the feasible bit; always
// records the lowest-index node whose score equals `highest`.
// ------------------------------------------------------------------------------------------
#define DCT 128  // nodes (threads) per CTA
template <bool WRITE>
__global__ void __launch_bounds__(DCT) k_direct_out(DevSnap s, const PodC *__restrict__ pods, u32 P, u32 pods_per_cta,
                                                    const u64 *__restrict__ max6, u64 *count,
                                                    const u64 *__restrict__ hilo, u64 *ckey, i64 *raw_out, u8 *norm_out,
                                                    u32 *bits_out, u64 stride, const u32 *gate, int gate_mode) {
    if (gate_closed(gate, gate_mode)) return;
    const u32 n = blockIdx.x * DCT + threadIdx.x;
    const bool live = n < s.n_nodes;
    u64 fr[G], tot[G], ck[G], co[G], bw[G], pw[G];
    u32 meta = 0;
    u64 cardnum = 0, stat = 0;
    if (live) {
        meta = s.meta[n];
        cardnum = s.card_number[n];
        stat = s.stat[n];
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const size_t o = (size_t)n * G + g;
            fr[g] = s.f[F_FREE][o];
            tot[g] = s.f[F_TOTAL][o];
            ck[g] = s.f[F_CLOCK][o];
            co[g] = s.f[F_CORE][o];
            bw[g] = s.f[F_BW][o];
            pw[g] = s.f[F_POWER][o];
        }
    } else {
#pragma unroll
        for (int g = 0; g < G; ++g) fr[g] = tot[g] = ck[g] = co[g] = bw[g] = pw[g] = 0;
    }
    const u32 p0 = blockIdx.y * pods_per_cta, p1 = min(P, p0 + pods_per_cta);
    for (u32 p = p0; p < p1; ++p) {
        const PodC pod = pods[p];
        u64 mx[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) mx[i] = max6[(size_t)p * 6 + i];
        i64 lo, hi;
        dev_lo_hi(hilo, p, lo, hi);
        u32 qual;
        const bool fit = direct_fit(pod, meta, cardnum, fr, ck, qual);
        u64 basic = 0;
        if (fit) {
#pragma unroll
            for (int g = 0; g < G; ++g)
                if ((qual >> g) & 1u) basic += direct_card_score(mx, bw[g], ck[g], co[g], pw[g], fr[g], tot[g]);
        }
        const i64 raw = (meta & META_HAS_SCV) ? dev_u64_to_i64(basic + stat) : 0;
        const u32 fitmask = __ballot_sync(0xFFFFFFFFu, fit);
        const u32 topmask = __ballot_sync(0xFFFFFFFFu, fit && raw == hi);
        if (topmask && (threadIdx.x & 31) == (u32)(__ffs(topmask) - 1)) atomicMax(&ckey[p], ckey_of(s.node_base + n));
        if (fitmask && (threadIdx.x & 31) == 0) atomicAdd(&count[p], (u64)__popc(fitmask));
        if (WRITE) {
            if (live) {
                raw_out[(size_t)p * stride + n] = raw;
                norm_out[(size_t)p * stride + n] = fit ? dev_norm_u8(dev_norm_i64(raw, lo, hi)) : (u8)0;
            }
            if ((threadIdx.x & 31) == 0 && (n - (threadIdx.x & 31)) < stride)
                bits_out[(size_t)p * (stride / 32) + (n >> 5)] = fitmask;
        }
    }
}

// Assemble yoda_pod_result from the (already cross-shard reduced) partial buffers.
// packed_key: ckey holds (top score << 32 | ~global index) instead of 2^32 - global index (merged DECIDE pass)
__global__ void k_finalize(const u64 *__restrict__ max6, const u64 *__restrict__ count, const u64 *__restrict__ hilo,
                           const u64 *__restrict__ ckey, yoda_pod_result *__restrict__ out, u32 P, bool packed_key) {
    u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    yoda_pod_result r;
    r.max_bandwidth = max6[(size_t)p * 6 + 0];
    r.max_clock = max6[(size_t)p * 6 + 1];
    r.max_core = max6[(size_t)p * 6 + 2];
    r.max_free_memory = max6[(size_t)p * 6 + 3];
    r.max_power = max6[(size_t)p * 6 + 4];
    r.max_total_memory = max6[(size_t)p * 6 + 5];
    i64 lo, hi;
    dev_lo_hi(hilo, p, lo, hi);
    r.lowest = lo;
    r.highest = hi;
    r.feasible = (u32)count[p];
    r.chosen = packed_key ? (count[p] ? (i32)(0xFFFFFFFFu - (u32)ckey[p]) : -1) : ckey_to_idx(ckey[p]);
    out[p] = r;
}

// ------------------------------------------------------------------------------------------
// Yoda.NormalizeScore on an arbitrary int64 list (scheduler.go:132-157)
// ------------------------------------------------------------------------------------------
__global__ void k_norm_list_reduce(const i64 *__restrict__ sc, u32 n, i64 *lohi /* [0]=lowest init scores[0], [1]=highest init 0 */) {
    i64 lo = INT64_MAX, hi = INT64_MIN;
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        lo = min(lo, sc[i]);
        hi = max(hi, sc[i]);
    }
    for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&lohi[0], lo);
        atomicMax(&lohi[1], hi);
    }
}
__global__ void k_norm_list_apply(i64 *sc, u32 n, const i64 *__restrict__ lohi) {
    i64 lo = lohi[0], hi = lohi[1];
    if (hi == lo) lo = (i64)((u64)lo - 1ull);
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) sc[i] = dev_norm_i64(sc[i], lo, hi);
}
