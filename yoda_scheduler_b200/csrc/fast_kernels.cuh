// fast_kernels.cuh -- the narrow (uint32) fast path.
//
// Applies when the derive step (prepare_node) found every card field and CardNumber < 2^32-1 and every static score
// < 2^24 (any real cluster).  It produces bit-identical results to the direct kernels with far
// fewer integer operations per (pod, node) pair:
//
//  1. pods are hashed into requirement groups (has-clock, clock, number, has-memory) and sorted by
//     (group, memory);
//  2. collection.CollectMaxValues becomes a 1-D dominance problem per group: a card counts for a
//     pod iff memory <= e(card) with e = min(FreeMemory, k-th largest healthy FreeMemory of its
//     node), so each card that no other card of its node dominates searches the group's sorted
//     memory values once and a suffix-max scan over the pods finishes the six maxima --
//     O((cards + pods) log pods) per group instead of O(pods x cards);
//  3. pods with the same six maxima form a class; per class a table of per-card scores
//     (algorithm.go:57-68, the six divisions taken with the class's reciprocals) is built once per batch;
//  4. the two hot kernels (score reduce, output pass) run lane = node with the node's eight
//     FreeMemory values and eight class card scores in registers and loop over the sorted pods:
//     one compare + one predicated add per card, the healthy-card count for Filter riding in
//     bits 24..27 of the same sum.
#pragma once
#include "common.cuh"
#include "direct_kernels.cuh"  // dev_lo_hi / dev_norm_i64: the exact rescale of pods outside the 32-bit range

// sorted pod record (32 B), broadcast to every lane of the hot kernels
struct PodS {
    u32 m, c, num;  // memory / clock / number, clamped to 2^32-1
    u32 req;        // bits 0-7 reqM (number if scv/memory present else 0), bits 8-15 reqC, bits 16-18 label flags
    u32 cls;        // class id (index of the card-score table)
    u32 orig;       // index in the caller's batch
    u32 gid;        // requirement group
    u32 pad;
};
struct GroupDesc {
    u32 flags, c, num, seg_start, seg_end, pad0, pad1, pad2;
};

#define CS_HEALTH_SHIFT 24
#define CS_SCORE_MASK 0xFFFFFFu
#define CS_LIMIT (1u << 21)  // admissible card scores must stay below this (8 cards -> sum < 2^24)

#define FAST_GID_BITS 12
#define FAST_GROUP_CAP (1u << FAST_GID_BITS)  // more requirement groups than this -> direct path
#define FAST_ERR_HASH 1u        // 64-bit hash collision between two different tuples
#define FAST_ERR_WIDE_CLASS 2u  // a class table entry does not fit the packed format
#define FAST_ERR_CAP 4u         // group/class capacity exceeded

// counters shared between the grouping kernels and the host
struct FastCtl {
    u32 n_groups, n_classes, reserved, err;
};

__device__ __forceinline__ u32 clamp32(u64 v) { return v > 0xFFFFFFFFull ? 0xFFFFFFFFu : (u32)v; }
__device__ __forceinline__ u64 mix64(u64 z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__device__ __forceinline__ void load8(const u32 *p, u32 *dst) {
    const uint4 a = reinterpret_cast<const uint4 *>(p)[0], b = reinterpret_cast<const uint4 *>(p)[1];
    dst[0] = a.x; dst[1] = a.y; dst[2] = a.z; dst[3] = a.w; dst[4] = b.x; dst[5] = b.y; dst[6] = b.z; dst[7] = b.w;
}

// ------------------------------------------------------------------------------------------
// open-addressing set of 64-bit keys; empty = all ones.  Insert only publishes the key; dense ids
// are handed out by k_hash_compact, so no thread ever waits on another.
// ------------------------------------------------------------------------------------------
#define HASH_EMPTY 0xFFFFFFFFFFFFFFFFull
__device__ __forceinline__ u32 hash_insert(u64 *keys, u32 mask, u64 key) {
    u32 slot = (u32)mix64(key) & mask;
    for (;;) {
        u64 cur = __ldcg(&keys[slot]);  // from L2: a key another SM has just published is seen, and no atomic is needed for it
        if (cur == key) return slot;
        if (cur == HASH_EMPTY) {
            cur = atomicCAS(&keys[slot], HASH_EMPTY, key);
            if (cur == HASH_EMPTY || cur == key) return slot;
        }
        slot = (slot + 1) & mask;
    }
}
__global__ void k_hash_clear(u64 *keys, u32 *rep, u32 n) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        keys[i] = HASH_EMPTY;
        rep[i] = 0xFFFFFFFFu;
    }
}
// occupied slot -> dense id (order arbitrary; results never depend on it)
__global__ void k_hash_compact(const u64 *keys, u32 *ids, u32 n, u32 *counter) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && keys[i] != HASH_EMPTY) ids[i] = atomicAdd(counter, 1u);
}

// ------------------------------------------------------------------------------------------
// step 1: requirement groups
// ------------------------------------------------------------------------------------------
// group key: everything Filter/CollectMaxValues read from the pod except the memory value
__device__ __forceinline__ u64 group_key(const PodC &p, u32 &num32, u32 &c32, u32 &gflags) {
    num32 = clamp32(p.num);
    c32 = (p.flags & YODA_POD_HAS_CLOCK) ? clamp32(p.clk) : 0u;
    gflags = p.flags & (YODA_POD_HAS_MEMORY | YODA_POD_HAS_CLOCK);
    if (gflags) {
        // with a per-card count requirement, every number > 8 is equally infeasible (<= 8 cards per node)
        if (num32 > G) { num32 = G + 1; c32 = 0; }
        return (1ull << 63) | ((u64)(gflags & YODA_POD_HAS_MEMORY ? 1 : 0) << 62) | ((u64)(gflags & YODA_POD_HAS_CLOCK ? 1 : 0) << 61) |
               ((u64)num32 << 32) | c32;
    }
    return (u64)num32;
}

// A batch has far fewer groups (and classes) than pods: a thread per pod would send the whole batch into compare-and-swaps
// on a hundred addresses at the same instant.  INSERT_PER pods per thread, one after the other: only the first round finds
// empty slots, the later ones read the published keys.
#define INSERT_PER 8
__global__ void k_group_insert(const PodC *__restrict__ pods, u32 P, u64 *gkeys, u32 gmask, u32 *slot_of_pod) {
    const u32 stride = gridDim.x * blockDim.x;
    for (u32 p = blockIdx.x * blockDim.x + threadIdx.x; p < P; p += stride) {
        u32 n32, c32, gf;
        const u64 key = group_key(pods[p], n32, c32, gf);
        slot_of_pod[p] = hash_insert(gkeys, gmask, key);
    }
}

// pod -> gid, sort key (has-clock | gid | memory), group descriptor
__global__ void k_group_assign(const PodC *__restrict__ pods, u32 P, const u32 *__restrict__ slot_of_pod,
                               const u32 *__restrict__ gids, GroupDesc *gdesc, u32 gcap, u64 *sort_keys, u32 *sort_vals,
                               FastCtl *ctl) {
    u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const PodC pod = pods[p];
    u32 n32, c32, gf;
    group_key(pod, n32, c32, gf);
    const u32 gid = gids[slot_of_pod[p]];
    const u32 m32 = clamp32(pod.mem);
    sort_keys[p] = ((u64)(gf & YODA_POD_HAS_CLOCK ? 1 : 0) << (32 + FAST_GID_BITS)) | ((u64)(gid & (FAST_GROUP_CAP - 1)) << 32) | m32;
    sort_vals[p] = p;
    if (gid < gcap) {
        // every pod of the group writes the same values
        gdesc[gid].flags = gf;
        gdesc[gid].c = c32;
        gdesc[gid].num = n32;
    } else {
        atomicOr(&ctl->err, FAST_ERR_CAP);
    }
}

// Small batches (<= FAST_SMALL_P pods, the one-pod-at-a-time scheduling cycle): every pod is its own
// group and its own class, so hashing, sorting and the host round trips that size their grids are
// skipped altogether.
#define FAST_SMALL_P 8
__global__ void k_small_groups(const PodC *__restrict__ pods, u32 P, GroupDesc *gdesc, PodS *out, u32 *m_sorted) {
    const u32 p = threadIdx.x;
    if (p >= P) return;
    const PodC pod = pods[p];
    u32 n32, c32, gf;
    group_key(pod, n32, c32, gf);
    GroupDesc g;
    g.flags = gf; g.c = c32; g.num = n32; g.seg_start = p; g.seg_end = p + 1; g.pad0 = g.pad1 = g.pad2 = 0;
    gdesc[p] = g;
    PodS s;
    s.m = clamp32(pod.mem);
    s.c = (pod.flags & YODA_POD_HAS_CLOCK) ? clamp32(pod.clk) : 0u;
    s.num = clamp32(pod.num);
    const u32 req = min(s.num, 255u);
    s.req = ((pod.flags & YODA_POD_HAS_MEMORY) ? req : 0u) | (((pod.flags & YODA_POD_HAS_CLOCK) ? req : 0u) << 8) | (pod.flags << 16);
    s.cls = p; s.orig = p; s.gid = p; s.pad = 0;
    out[p] = s;
    m_sorted[p] = s.m;
}
// The whole set-up of a small batch in one launch (a one-pod cycle is launch-bound): what k_pod_prep, k_init_partials,
// k_small_groups and the clearing of the step planes and of the control block do for bigger batches.
__global__ void k_small_setup(const yoda_pod_req *__restrict__ req, PodC *pods, u32 P, u64 *max6, u64 *count, u64 *hilo, u64 *ckey, GroupDesc *gdesc,
                              PodS *out, u32 *m_sorted, u32 *step, FastCtl *ctl) {
    const u32 t = threadIdx.x;
    if (t < P * 6) {
        step[t] = 0;
        max6[t] = 1ull;  // collection.go:31-38
    }
    if (t == 0) {
        ctl->n_groups = 0; ctl->n_classes = 0; ctl->reserved = 0; ctl->err = 0;
    }
    if (t >= P) return;
    const yoda_pod_req r = req[t];
    PodC pod;
    pod.flags = r.flags & 7u;
    pod.num = (r.flags & YODA_POD_HAS_NUMBER) ? r.number : 1ull;  // filter.go:11-16
    pod.mem = (r.flags & YODA_POD_HAS_MEMORY) ? r.memory : 0ull;  // filter.go:18-33
    pod.clk = (r.flags & YODA_POD_HAS_CLOCK) ? r.clock : 0ull;    // filter.go:35-50
    pod.pad = 0;
    pods[t] = pod;
    count[t] = 0; hilo[2 * t] = 0; hilo[2 * t + 1] = 0; ckey[t] = 0;
    u32 n32, c32, gf;
    group_key(pod, n32, c32, gf);
    GroupDesc g;
    g.flags = gf; g.c = c32; g.num = n32; g.seg_start = t; g.seg_end = t + 1; g.pad0 = g.pad1 = g.pad2 = 0;
    gdesc[t] = g;
    PodS ps;
    ps.m = clamp32(pod.mem);
    ps.c = (pod.flags & YODA_POD_HAS_CLOCK) ? clamp32(pod.clk) : 0u;
    ps.num = clamp32(pod.num);
    const u32 rq = min(ps.num, 255u);
    ps.req = ((pod.flags & YODA_POD_HAS_MEMORY) ? rq : 0u) | (((pod.flags & YODA_POD_HAS_CLOCK) ? rq : 0u) << 8) | (pod.flags << 16);
    ps.cls = t; ps.orig = t; ps.gid = t; ps.pad = 0;
    out[t] = ps;
    m_sorted[t] = ps.m;
}

__global__ void k_small_classes(u32 P, const u64 *__restrict__ max6, u32 *cls_max, FastCtl *ctl) {
    const u32 t = threadIdx.x;
    if (t >= P * 6) return;
    const u64 v = max6[t];
    if (v > 0xFFFFFFFFull) atomicOr(&ctl->err, FAST_ERR_WIDE_CLASS);  // another shard may be wide
    cls_max[t] = (u32)v;
}

// sorted pod records + group segment bounds
__global__ void k_build_sorted(const PodC *__restrict__ pods, u32 P, const u64 *__restrict__ keys_sorted,
                               const u32 *__restrict__ vals_sorted, GroupDesc *gdesc, u32 gcap, PodS *out, u32 *m_sorted) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const u64 key = keys_sorted[i];
    const u32 orig = vals_sorted[i];
    const PodC pod = pods[orig];
    u32 n32, c32, gf;
    group_key(pod, n32, c32, gf);
    const u32 gid = (u32)(key >> 32) & (FAST_GROUP_CAP - 1);
    PodS s;
    s.m = (u32)key;
    s.c = (pod.flags & YODA_POD_HAS_CLOCK) ? clamp32(pod.clk) : 0u;
    s.num = clamp32(pod.num);
    const u32 req = min(s.num, 255u);
    s.req = ((pod.flags & YODA_POD_HAS_MEMORY) ? req : 0u) | (((pod.flags & YODA_POD_HAS_CLOCK) ? req : 0u) << 8) | (pod.flags << 16);
    s.cls = 0;
    s.orig = orig;
    s.gid = gid;
    s.pad = 0;
    out[i] = s;
    m_sorted[i] = s.m;
    const bool head = (i == 0) || ((u32)(keys_sorted[i - 1] >> 32) != (u32)(key >> 32));
    const bool tail = (i == P - 1) || ((u32)(keys_sorted[i + 1] >> 32) != (u32)(key >> 32));
    if (gid < gcap) {
        if (head) gdesc[gid].seg_start = i;
        if (tail) gdesc[gid].seg_end = i + 1;
    }
}

// ------------------------------------------------------------------------------------------
// step 2: CollectMaxValues as dominance queries.  thread = (node, group).
// step[f][i] (sorted pod position i) receives the maxima of the cards whose LAST admissible pod is i;
// k_max_tile_lead + k_max_scan turn that into the maxima of all cards admissible for pod i.
// ------------------------------------------------------------------------------------------
#define MAXT 128
#define MAXG_TILE 8  // group descriptors staged per shared-memory tile
// The group count is taken from the device (ctl) when the host has not read it back (one-GPU batches
// run without that round trip); grids are then sized for the worst case and surplus blocks leave at once.
__device__ __forceinline__ u32 group_count(const FastCtl *ctl, u32 n_groups) { return ctl ? min(ctl->n_groups, FAST_GROUP_CAP) : n_groups; }

__global__ void __launch_bounds__(MAXT) k_max_thresholds(DevSnap s, const GroupDesc *__restrict__ gdesc, u32 n_groups_host,
                                                         const FastCtl *__restrict__ ctl, const u32 *__restrict__ m_sorted, u32 *step, u32 P) {
    __shared__ GroupDesc sg[MAXG_TILE];
    const u32 n_groups = group_count(ctl, n_groups_host);
    if (blockIdx.y * MAXG_TILE >= n_groups) return;
    const u32 slot = blockIdx.x * MAXT + threadIdx.x;
    const bool live = slot < s.n_nodes;
    const u32 nn = live ? s.order[slot] : 0;  // nodes with the same clock sit in the same warps
    const u32 meta = live ? s.meta[nn] : 0u;
    const bool has = meta_table_node(meta);  // wide nodes: the exact kernels add their maxima afterwards
    const u32 hm = meta_healthy(meta), keep = meta_keep(meta);
    const u32 cardnum = s.cardnum32[nn];
    const u32 n_healthy = __popc(hm);
    // Tk[k-1] = k-th largest FreeMemory among healthy cards: the node admits memory values m <= Tk[num-1]
    // when scv/memory is present (filter.go:18-33)
    u32 ck[G], Tk[G];
    load8(s.n32[F_CLOCK] + (size_t)nn * G, ck);
    load8(s.tk32 + (size_t)nn * G, Tk);
    // blockIdx.y takes every gridDim.y-th tile of groups
    for (u32 g0 = blockIdx.y * MAXG_TILE; g0 < n_groups; g0 += MAXG_TILE * gridDim.y) {
        __syncthreads();
        if (threadIdx.x < MAXG_TILE && g0 + threadIdx.x < n_groups) sg[threadIdx.x] = gdesc[g0 + threadIdx.x];
        __syncthreads();
        const u32 gn = min(MAXG_TILE, n_groups - g0);
        if (!has) continue;
        // which groups of the tile this node is feasible for, and the memory bound T it admits in each
        u32 T[MAXG_TILE], pass = 0;
#pragma unroll
        for (int gi = 0; gi < MAXG_TILE; ++gi) {
            T[gi] = 0xFFFFFFFFu;
            if (gi >= (int)gn) continue;
            const GroupDesc gd = sg[gi];
            bool ok = gd.num <= cardnum && gd.seg_end > gd.seg_start;  // filter.go:13,15
            if (gd.flags & YODA_POD_HAS_CLOCK) {                       // filter.go:35-50
                u32 cnt = 0;
#pragma unroll
                for (int g = 0; g < G; ++g) cnt += ((hm >> g) & 1u) & (u32)(ck[g] == gd.c);
                ok = ok && cnt >= gd.num;
            }
            if ((gd.flags & YODA_POD_HAS_MEMORY) && gd.num >= 1) {     // filter.go:18-33
                ok = ok && n_healthy >= gd.num;
                u32 t = 0;
#pragma unroll
                for (int k = 0; k < G; ++k)
                    if (gd.num == (u32)(k + 1)) t = Tk[k];
                T[gi] = t;
            }
            pass |= ok ? (1u << gi) : 0u;
        }
        if (!pass) continue;
        // Only cards that no other card of the node dominates can hold a maximum (a dominating card is admitted
        // by every pod that admits the dominated one); homogeneous nodes keep a single card.
        for (u32 km = keep; km; km &= km - 1) {
            const int g = __ffs(km) - 1;
            const size_t o = (size_t)nn * G + g;
            const u32 frg = s.n32[F_FREE][o], ckg = s.n32[F_CLOCK][o];
            // j[gi] = number of pods of group gi with memory <= min(frg, T[gi]): branch-free upper bounds, the
            // groups of the tile side by side so that their loads are in flight together
            u32 j[MAXG_TILE], rem[MAXG_TILE], more = 0;
#pragma unroll
            for (int gi = 0; gi < MAXG_TILE; ++gi) {
                j[gi] = 0;
                const bool act = ((pass >> gi) & 1u) && ckg >= sg[gi].c;  // collection.go:46 (clock >= c; c == 0 without the label)
                rem[gi] = act ? sg[gi].seg_end - sg[gi].seg_start : 0u;
                more |= rem[gi] > 1 ? 1u : 0u;
            }
            while (more) {
                more = 0;
#pragma unroll
                for (int gi = 0; gi < MAXG_TILE; ++gi) {
                    if (rem[gi] > 1) {
                        const u32 half = rem[gi] >> 1;
                        j[gi] += (m_sorted[sg[gi].seg_start + j[gi] + half - 1] <= min(frg, T[gi])) ? half : 0u;
                        rem[gi] -= half;
                        more |= rem[gi] > 1 ? 1u : 0u;
                    }
                }
            }
            // collection.go:59-78; order of the step planes = collection.MaxValue (bw, clock, core, free, power, total)
            const u32 v[6] = {s.n32[F_BW][o], ckg, s.n32[F_CORE][o], frg, s.n32[F_POWER][o], s.n32[F_TOTAL][o]};
#pragma unroll
            for (int gi = 0; gi < MAXG_TILE; ++gi) {
                if (!rem[gi]) continue;
                const u32 cnt = j[gi] + ((m_sorted[sg[gi].seg_start + j[gi]] <= min(frg, T[gi])) ? 1u : 0u);
                if (!cnt) continue;
                const u32 pos = sg[gi].seg_start + cnt - 1;
                // Short segments (small batches above all) send every lane to the same position: one atomic per warp
                // then, not one per lane.  Any set of lanes that happens to be here together may team up.
                const u32 team = __activemask();
                const u32 lead = __ffs(team) - 1;
                const bool same = __all_sync(team, pos == __shfl_sync(team, pos, lead));
#pragma unroll
                for (int f = 0; f < 6; ++f) {
                    u32 *dst = &step[(size_t)f * P + pos];
                    if (same) {
                        const u32 m = __reduce_max_sync(team, v[f]);
                        if ((threadIdx.x & 31) == lead && m > __ldcg(dst)) atomicMax(dst, m);
                    } else if (v[f] > __ldcg(dst)) {  // L2: sees the other SMs' maxima
                        atomicMax(dst, v[f]);
                    }
                }
            }
        }
    }
}

// Suffix maxima of the step planes over every group's pod segment, result max(1, .) into max6[orig].
// Positions are cut into tiles of SCANT regardless of the groups, so one huge group costs no more than
// many small ones.  k_max_tile_lead: per (tile, field) the maximum over the tile's leading run of positions
// that belong to the group of its first position -- what the tiles before it need.  k_max_scan: a segmented
// suffix scan inside the tile (groups are contiguous; a position's run ends at its group's seg_end), plus,
// for the group of the tile's last position, the leads of the later tiles that start inside that group.
#define SCANT 1024
__global__ void __launch_bounds__(SCANT) k_max_tile_lead(const GroupDesc *__restrict__ gdesc, const u32 *__restrict__ step,
                                                         const PodS *__restrict__ pods, u32 *lead, u32 P, u32 n_tiles) {
    __shared__ u32 warp_max[SCANT / 32];
    const u32 tile = blockIdx.x, f = blockIdx.y;
    const u32 t0 = tile * SCANT, pos = t0 + threadIdx.x;
    const u32 run_end = min(min(P, t0 + SCANT), gdesc[pods[t0].gid].seg_end);
    u32 v = pos < run_end ? step[(size_t)f * P + pos] : 0u;
    v = __reduce_max_sync(0xFFFFFFFFu, v);
    if ((threadIdx.x & 31) == 0) warp_max[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = __reduce_max_sync(0xFFFFFFFFu, warp_max[threadIdx.x]);
        if (threadIdx.x == 0) lead[(size_t)f * n_tiles + tile] = v;
    }
}

__global__ void __launch_bounds__(SCANT) k_max_scan(const GroupDesc *__restrict__ gdesc, const u32 *__restrict__ step,
                                                    const PodS *__restrict__ pods, const u32 *__restrict__ lead, u64 *max6, u32 P,
                                                    u32 n_tiles) {
    __shared__ u32 warp_lead[SCANT / 32];  // per warp: the run of its first position, as far as it goes inside the warp
    __shared__ u32 red[SCANT / 32];
    const u32 tile = blockIdx.x, f = blockIdx.y;
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u32 t0 = tile * SCANT, t1 = min(P, t0 + SCANT), pos = t0 + threadIdx.x;
    const bool live = pos < t1;
    const uint4 hi = live ? reinterpret_cast<const uint4 *>(pods + pos)[1] : make_uint4(0, 0, 0xFFFFFFFFu, 0);  // cls, orig, gid, pad
    const u32 gid = hi.z, orig = hi.y;
    u32 v = live ? step[(size_t)f * P + pos] : 0u;
    // inside the warp: suffix maximum over the lanes after this one that are in the same group
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u32 tv = __shfl_down_sync(0xFFFFFFFFu, v, o), tg = __shfl_down_sync(0xFFFFFFFFu, gid, o);
        if (lane + o < 32 && tg == gid) v = max(v, tv);
    }
    if (lane == 0) warp_lead[warp] = v;
    // the group of the tile's last position may go on in later tiles
    const u32 g_tail = pods[t1 - 1].gid;
    const u32 tail_end = gdesc[g_tail].seg_end;
    u32 carry = 0;
    for (u32 t = tile + 1 + threadIdx.x; t < n_tiles && t * SCANT < tail_end; t += SCANT) carry = max(carry, lead[(size_t)f * n_tiles + t]);
    carry = __reduce_max_sync(0xFFFFFFFFu, carry);
    if (lane == 0) red[warp] = carry;
    __syncthreads();
    // later warps of this tile that start inside the group of this warp's last lane
    const u32 g_last = __shfl_sync(0xFFFFFFFFu, gid, 31);
    const u32 last_end = g_last == 0xFFFFFFFFu ? 0u : gdesc[g_last].seg_end;
    u32 pre = (lane > warp && t0 + lane * 32 < min(last_end, t1)) ? warp_lead[lane] : 0u;  // SCANT / 32 == 32 warps: one per lane
    pre = __reduce_max_sync(0xFFFFFFFFu, pre);
    carry = __reduce_max_sync(0xFFFFFFFFu, red[lane]);
    if (gid == g_last) v = max(v, pre);
    if (gid == g_tail) v = max(v, carry);
    if (live) max6[(size_t)orig * 6 + f] = (u64)max(v, 1u);  // collection.go:31-38: the maxima start at 1
}

// small batches (k_small_groups): every pod is its own group, so its maxima are its own step entries
__global__ void k_small_max_out(const u32 *__restrict__ step, const PodS *__restrict__ pods, u64 *max6, u32 P) {
    const u32 i = threadIdx.x / 6, f = threadIdx.x % 6;
    if (i < P) max6[(size_t)pods[i].orig * 6 + f] = (u64)max(step[(size_t)f * P + i], 1u);  // collection.go:31-38
}

// ------------------------------------------------------------------------------------------
// step 3: classes = distinct (reduced) maxima tuples
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 class_hash(const u64 *mx) {
    u64 h = 0x9E3779B97F4A7C15ull;
#pragma unroll
    for (int i = 0; i < 6; ++i) h = mix64(h ^ mx[i]) + 0x9E3779B97F4A7C15ull * (i + 1);
    return h & 0x7FFFFFFFFFFFFFFFull;  // never HASH_EMPTY
}
__global__ void k_class_insert(const PodS *__restrict__ pods, u32 P, const u64 *__restrict__ max6, u64 *ckeys, u32 cmask, u32 *rep,
                               u32 *slot_of_pod) {
    // (a thread per pod here: the six maxima of a pod are a scattered 48-byte read, which wants all the threads it can get)
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const u32 orig = pods[i].orig;
    const u32 slot = hash_insert(ckeys, cmask, class_hash(max6 + (size_t)orig * 6));
    slot_of_pod[i] = slot;
    if (__ldcg(&rep[slot]) > orig) atomicMin(&rep[slot], orig);  // a deterministic representative pod
}
// pod -> class id; verify the tuple against the representative's (hash collisions -> error flag);
// the representative records the class maxima
__global__ void k_class_assign(PodS *pods, u32 P, const u64 *__restrict__ max6, const u32 *__restrict__ rep,
                               const u32 *__restrict__ cids, const u32 *__restrict__ slot_of_pod, u32 *cls_max, u32 ccap, FastCtl *ctl) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const u32 slot = slot_of_pod[i];
    const u32 cid = cids[slot], r = rep[slot], orig = pods[i].orig;
    bool same = true, narrow = true;
#pragma unroll
    for (int f = 0; f < 6; ++f) {
        same &= max6[(size_t)orig * 6 + f] == max6[(size_t)r * 6 + f];
        narrow &= max6[(size_t)orig * 6 + f] <= 0xFFFFFFFFull;  // another shard may be wide
    }
    if (!same) atomicOr(&ctl->err, FAST_ERR_HASH);
    if (!narrow) atomicOr(&ctl->err, FAST_ERR_WIDE_CLASS);
    pods[i].cls = cid < ccap ? cid : 0u;  // (a batch over the capacity is flagged and redone; until then no table index may leave the buffer)
    if (cid >= ccap) {
        atomicOr(&ctl->err, FAST_ERR_CAP);
    } else if (orig == r) {
#pragma unroll
        for (int f = 0; f < 6; ++f) cls_max[(size_t)cid * 6 + f] = (u32)max6[(size_t)r * 6 + f];
    }
}

// Node-sharded batches that ran with the counts on the device: 1 if this shard hit a capacity limit (the batch has to
// be redone the careful way); all-reduced with MAX so that every shard takes the same decision.
__global__ void k_cap_flag(const FastCtl *__restrict__ ctl, u64 *flag) { *flag = (ctl->err & FAST_ERR_CAP) ? 1ull : 0ull; }

// Per class, per card: CalculateCardScore (algorithm.go:57-68) packed with the card's health bit.
// A card whose fields exceed the class maxima can never be scored by a pod of that class (it would
// have raised the maxima), so its score slot is 0; its health bit stays (Filter still counts it).
//
// floor(x * 100 / d) is taken with the divisor's reciprocal M = ceil(2^64 / d), d >= 2: for every
// a = 100x < 2^32, mulhi64(a, M) == floor(a / d), because M*d = 2^64 + e with 0 <= e < d and a*e < 2^64.
// With a in 32 bits that is one 32x32 high product and one 32x32+64 multiply-add.  Nodes with a field
// >= 2^32/100 and classes with a maximum of 1 (no 64-bit reciprocal) take the plain-division path.
#define CLST 128
#define CLS_PER 8  // classes per thread: the node's 48 field values are loaded once, and quotients are reused between classes

__device__ __forceinline__ u32 rdiv(u32 a, u32 Mhi, u32 Mlo) {
    return (u32)(((u64)a * Mhi + __umulhi(a, Mlo)) >> 32);  // a*Mhi + carry < 2^64
}

// plain 64-bit arithmetic for one card; dm = the class maxima in cls_max order (bw, clock, core, free, power, total)
// bit 31 of the result = "a term left the packed range"
__device__ __forceinline__ u32 card_score_plain(const DevSnap &s, u32 n, int g, const u32 *dm) {
    const size_t o = (size_t)n * G + g;
    const u32 bw = s.n32[F_BW][o], ck = s.n32[F_CLOCK][o], co = s.n32[F_CORE][o], pw = s.n32[F_POWER][o], fr = s.n32[F_FREE][o],
              to = s.n32[F_TOTAL][o];
    if (!(bw <= dm[0] && ck <= dm[1] && co <= dm[2] && fr <= dm[3] && pw <= dm[4] && to <= dm[5])) return 0;
    const u64 q0 = (u64)bw * 100ull / dm[0], q1 = (u64)ck * 100ull / dm[0], q2 = (u64)co * 100ull / dm[2], q3 = (u64)pw * 100ull / dm[4],
              q4 = (u64)fr * 100ull / dm[3], q5 = (u64)to * 100ull / dm[5];
    const u64 cs = q0 + q1 + q2 + q3 + q4 * 2ull + q5;  // every term < 2^39
    return ((u32)cs & CS_SCORE_MASK) | (((q0 | q1 | q2 | q3 | q4 | q5 | cs) >= CS_LIMIT) ? 0x80000000u : 0u);
}

// Classes in the order the table kernel walks them: sorted by (bandwidth, core, power, total, clock, free) maxima, so that
// classes of one "family" -- the same best card model, i.e. equal maxima except for MaxFreeMemory, which is what
// distinguishes most classes of a real batch -- sit next to each other and share five of their six quotients.
// Rank by counting (n <= 4096 classes, usually < 100); order[rank] = class id.
// One warp per class, the lanes stride over the other classes.
#define CLS_ORDER_T 128
__global__ void __launch_bounds__(CLS_ORDER_T) k_class_order(const u32 *__restrict__ cls_max, u32 n_classes, const FastCtl *__restrict__ ctl,
                                                             bool count_on_device, u32 *order) {
    if (count_on_device) n_classes = min(n_classes, ctl->n_classes);
    const u32 i = (blockIdx.x * CLS_ORDER_T + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= n_classes) return;  // warp-uniform
    constexpr int key[6] = {0, 2, 4, 5, 1, 3};  // cls_max order is bw, clock, core, free, power, total
    u32 mine[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) mine[k] = cls_max[i * 6 + key[k]];
    u32 rank = 0;
    for (u32 j = lane; j < n_classes; j += 32) {
        bool lt = false, eq = true;  // class j before class i / equal so far
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const u32 o = cls_max[j * 6 + key[k]];
            lt = lt || (eq && o < mine[k]);
            eq = eq && o == mine[k];
        }
        rank += (lt || (eq && j < i)) ? 1u : 0u;  // the tuples are distinct; the index settles whatever is left
    }
    rank = __reduce_add_sync(0xFFFFFFFFu, rank);
    if (lane == 0) order[rank] = i;
}

struct ClassConsts {
    u32 d[6];     // maxima in cls_max order
    u32 d100[6];  // min(100 * d, 2^32 - 1): a <= d100 <=> x <= d for every a = 100x < 2^32
    u32 mhi[6], mlo[6];
    u32 plain;    // some maximum is 1
    u32 cid;      // the class this slot of the block computes
    u32 chg;      // bit f: maximum f differs from the previous reciprocal-path class of the block (all set for the first)
};
#define CLS_CHG_FAMILY 0x35u  // bandwidth, core, power, total: the divisors of five of the six quotients
#define CLS_CHG_CLOCK 0x02u   // MaxClock only gates (algorithm.go:60 divides the clock by MaxBandwidth)
#define CLS_CHG_FREE 0x08u

// order: the walk order (k_class_order), nullptr = class ids as they are (small batches: a class per pod).
// count_on_device: the class count is taken from ctl (the host sized the grid for the table capacity instead of
// reading the count back); surplus blocks leave at once.
__global__ void __launch_bounds__(CLST) k_class_tables(DevSnap s, const u32 *__restrict__ cls_max, u32 n_classes, u32 *tables,
                                                       FastCtl *ctl, bool count_on_device, const u32 *__restrict__ order) {
    __shared__ ClassConsts cc[CLS_PER];
    if (count_on_device) n_classes = min(n_classes, ctl->n_classes);
    const u32 pos0 = blockIdx.y * CLS_PER;
    if (pos0 >= n_classes) return;
    const u32 ncls = min((u32)CLS_PER, n_classes - pos0);
    if (threadIdx.x < CLS_PER) {
        cc[threadIdx.x].plain = 0;
        cc[threadIdx.x].cid = threadIdx.x < ncls ? (order ? order[pos0 + threadIdx.x] : pos0 + threadIdx.x) : 0u;
    }
    __syncthreads();
    if (threadIdx.x < ncls * 6) {
        const u32 k = threadIdx.x / 6, f = threadIdx.x % 6;
        const u32 d = cls_max[cc[k].cid * 6 + f];
        const u64 M = d > 1 ? (0xFFFFFFFFFFFFFFFFull / d) + 1ull : 0ull;  // ceil(2^64 / d), also for powers of two
        cc[k].d[f] = d;
        cc[k].d100[f] = (u32)min((u64)d * 100ull, 0xFFFFFFFFull);
        cc[k].mhi[f] = (u32)(M >> 32);
        cc[k].mlo[f] = (u32)M;
        if (d <= 1) atomicOr(&cc[k].plain, 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {  // which maxima change from one reciprocal-path class of the block to the next
        int prev = -1;
        for (u32 k = 0; k < ncls; ++k) {
            u32 chg = 0x3Fu;
            if (prev >= 0) {
                chg = 0;
                for (int f = 0; f < 6; ++f) chg |= (cc[k].d[f] != cc[prev].d[f]) ? (1u << f) : 0u;
            }
            cc[k].chg = chg;
            if (!cc[k].plain) prev = (int)k;
        }
    }
    __syncthreads();
    const u32 n = blockIdx.x * CLST + threadIdx.x;
    if (n >= s.n_nodes) return;
    const u32 meta = s.meta[n];
    if (meta & META_WIDE) {  // scored by the exact kernels; its narrow planes are truncations, keep them out of the tables
        for (u32 k = 0; k < ncls; ++k) {
            uint4 *dst = reinterpret_cast<uint4 *>(tables + ((size_t)cc[k].cid * s.cap + n) * G);
            dst[0] = make_uint4(0, 0, 0, 0);
            dst[1] = make_uint4(0, 0, 0, 0);
        }
        return;
    }
    const u32 hm = meta_healthy(meta);
    // a node's eight values of one field are 32 contiguous bytes: two 16-byte loads per field; kept as 100x
    u32 a[6][G];
    bool node_plain = false;
#pragma unroll
    for (int f = 0; f < 6; ++f) {
        const uint4 *src = reinterpret_cast<const uint4 *>(s.n32[f] + (size_t)n * G);
        const uint4 v0 = src[0], v1 = src[1];
        const u32 v[G] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int g = 0; g < G; ++g) {
            node_plain |= v[g] > 0xFFFFFFFFu / 100u;
            a[f][g] = v[g] * 100u;
        }
    }
    bool big = false;
    // carried from class to class (a thread on the plain path never reads them):
    u32 fam_sum[G], fam_or[G], q_free[G];  // bw + clock + core + power + total quotients, their OR, the FreeMemory quotient
    u32 adm_fam = 0, adm_clock = 0, adm_free = 0;  // bit g: card g within the class maxima of those fields
    for (u32 k = 0; k < ncls; ++k) {
        const ClassConsts &c = cc[k];
        u32 out[G];
        if (node_plain || c.plain) {
#pragma unroll 1
            for (int g = 0; g < G; ++g) {
                const u32 cs = card_score_plain(s, n, g, c.d);
                big |= (cs >> 31) != 0;
                const u32 w = (cs & CS_SCORE_MASK) | (((hm >> g) & 1u) << CS_HEALTH_SHIFT);
#pragma unroll
                for (int h = 0; h < G; ++h)
                    if (h == g) out[h] = w;
            }
        } else {
            const u32 chg = c.chg;  // block-uniform
            if (chg & CLS_CHG_FAMILY) {
                adm_fam = 0;
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    const u32 q0 = rdiv(a[F_BW][g], c.mhi[0], c.mlo[0]), q1 = rdiv(a[F_CLOCK][g], c.mhi[0], c.mlo[0]),
                              q2 = rdiv(a[F_CORE][g], c.mhi[2], c.mlo[2]), q3 = rdiv(a[F_POWER][g], c.mhi[4], c.mlo[4]),
                              q5 = rdiv(a[F_TOTAL][g], c.mhi[5], c.mlo[5]);
                    fam_sum[g] = q0 + q1 + q2 + q3 + q5;
                    fam_or[g] = q0 | q1 | q2 | q3 | q5;
                    const bool in = a[F_BW][g] <= c.d100[0] && a[F_CORE][g] <= c.d100[2] && a[F_POWER][g] <= c.d100[4] && a[F_TOTAL][g] <= c.d100[5];
                    adm_fam |= in ? (1u << g) : 0u;
                }
            }
            if (chg & CLS_CHG_CLOCK) {
                adm_clock = 0;
#pragma unroll
                for (int g = 0; g < G; ++g) adm_clock |= (a[F_CLOCK][g] <= c.d100[1]) ? (1u << g) : 0u;
            }
            if (chg & CLS_CHG_FREE) {
                adm_free = 0;
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    q_free[g] = rdiv(a[F_FREE][g], c.mhi[3], c.mlo[3]);
                    adm_free |= (a[F_FREE][g] <= c.d100[3]) ? (1u << g) : 0u;
                }
            }
            const u32 adm = adm_fam & adm_clock & adm_free;
#pragma unroll
            for (int g = 0; g < G; ++g) {
                const u32 cs = fam_sum[g] + q_free[g] * 2u;  // exact when every term < CS_LIMIT, flagged otherwise
                const bool ad = (adm >> g) & 1u;
                big |= ad && ((fam_or[g] | q_free[g] | cs) >= CS_LIMIT);
                out[g] = (ad ? (cs & CS_SCORE_MASK) : 0u) | (((hm >> g) & 1u) << CS_HEALTH_SHIFT);
            }
        }
        uint4 *dst = reinterpret_cast<uint4 *>(tables + ((size_t)c.cid * s.cap + n) * G);
        dst[0] = make_uint4(out[0], out[1], out[2], out[3]);
        dst[1] = make_uint4(out[4], out[5], out[6], out[7]);
    }
    if (big) atomicOr(&ctl->err, FAST_ERR_WIDE_CLASS);
}

// ------------------------------------------------------------------------------------------
// step 4: the hot kernels.  lane = node (NPL nodes per lane, interleaved by 32), loop over pods.
// ------------------------------------------------------------------------------------------
#define HOT_THREADS 256
#define HOT_WARPS (HOT_THREADS / 32)
#define HOT_CHUNK 128  // pods per CTA of the output pass
#ifndef SCORE_CHUNK
#define SCORE_CHUNK 256  // pods per CTA of the score pass: every chunk starts with a full evaluation of the CTA's nodes
                         // (measured best of 160/192/256/320; ~43 KB of shared memory per CTA, four CTAs per SM)
#endif

template <int NPL>
struct NodeRegs {
    u32 fr[NPL][G];   // FreeMemory of the node's card slots
    u32 cs[NPL][G];   // card contributions for the current (group, class): score bits | health bit << 24
    u32 cardnum[NPL];
    u32 stat[NPL];
    bool ok[NPL];     // node exists and has an SCV
    bool clkok[NPL];  // enough healthy cards with Clock == c for the current group (true without scv/clock)
};

template <int NPL>
__device__ __forceinline__ void hot_load_nodes(NodeRegs<NPL> &r, const DevSnap &s, u32 base, u32 lane) {
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
        const u32 n = base + (j >> 1) * 64 + 2 * lane + (j & 1);  // a lane owns pairs of adjacent nodes: 16-byte row pieces
        const bool in = n < s.n_nodes;
        const u32 nn = in ? n : 0;
        load8(s.n32[F_FREE] + (size_t)nn * G, r.fr[j]);
        r.cardnum[j] = s.cardnum32[nn];
        r.stat[j] = in ? s.stat32[nn] : 0u;
        r.ok[j] = in && meta_table_node(s.meta[nn]);
        r.clkok[j] = true;
    }
}

// (Re)load the per-card contributions when the class or the requirement group changes.  For a
// group with scv/clock the clock tests are folded in here, once per group instead of once per pod:
//   - Filter's clock count (healthy cards with Clock == c, filter.go:56-58) becomes the flag clkok;
//   - cards with Clock < c lose their score bits (collection.go:46 / algorithm.go:48 need
//     Clock >= c) but keep the health bit, because Filter's memory count ignores the clock.
template <int NPL>
__device__ __forceinline__ void hot_load_class(NodeRegs<NPL> &r, const DevSnap &s, const u32 *tables, u32 cls, bool has_clock, u32 c,
                                               u32 reqC, u32 base, u32 lane) {
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
        const u32 n = base + (j >> 1) * 64 + 2 * lane + (j & 1);
        const u32 nn = n < s.n_nodes ? n : 0;
        load8(tables + ((size_t)cls * s.cap + nn) * G, r.cs[j]);
        r.clkok[j] = true;
        if (has_clock) {  // warp-uniform
            u32 cnt = 0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {  // four cards at a time keeps the transient registers low
                const uint4 k4 = reinterpret_cast<const uint4 *>(s.n32[F_CLOCK] + (size_t)nn * G)[h];
                const u32 ck[4] = {k4.x, k4.y, k4.z, k4.w};
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    cnt += (ck[g] == c) ? (r.cs[j][4 * h + g] >> CS_HEALTH_SHIFT) : 0u;
                    if (ck[g] < c) r.cs[j][4 * h + g] &= ~CS_SCORE_MASK;
                }
            }
            r.clkok[j] = cnt >= reqC;
        }
    }
}

// Filter verdict and Score of (pod, node j): `fit`, raw = basic (if fit) + static part, and
// `bound` = the smallest FreeMemory of the node that is >= m.  For any m' in [m, bound] the set of
// cards with FreeMemory >= m' is the same, hence so are fit and raw (same group and class).
template <int NPL>
__device__ __forceinline__ u32 hot_eval(const NodeRegs<NPL> &r, int j, u32 m, u32 num, u32 reqM, bool &fit, u32 &bound) {
    u32 acc = 0, b = 0xFFFFFFFFu;
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const bool q = r.fr[j][g] >= m;  // filter.go:53 / collection.go:46 / algorithm.go:48
        b = q ? min(b, r.fr[j][g]) : b;
        acc += q ? r.cs[j][g] : 0u;
    }
    bound = b;
    const u32 cnt = acc >> CS_HEALTH_SHIFT;   // healthy cards with FreeMemory >= m (filter.go:52-54)
    const u32 basic = acc & CS_SCORE_MASK;    // algorithm.go:41-55
    fit = r.ok[j] && r.clkok[j] && (num <= r.cardnum[j]) && (cnt >= reqM);
    return r.stat[j] + (fit ? basic : 0u);    // algorithm.go:38 (no uint64 wrap in the narrow range)
}

// The same evaluation with the node's cards fetched on demand (L2) instead of held in registers:
// used by the score pass, where re-evaluations are rare and a small register footprint (more
// resident warps) matters more than their latency.
__device__ __forceinline__ u32 hot_eval_mem(const DevSnap &s, const u32 *__restrict__ tables, u32 node, u32 cls, u32 m, u32 num, u32 reqM,
                                            bool has_clock, u32 c, u32 reqC, bool ok, u32 cardnum, u32 stat, bool &fit, u32 &bound) {
    u32 fr[G], cs[G];
    load8(s.n32[F_FREE] + (size_t)node * G, fr);
    load8(tables + ((size_t)cls * s.cap + node) * G, cs);
    bool clkok = true;
    if (has_clock) {  // warp-uniform; see hot_load_class
        u32 ck[G];
        load8(s.n32[F_CLOCK] + (size_t)node * G, ck);
        u32 cnt = 0;
#pragma unroll
        for (int g = 0; g < G; ++g) {
            cnt += (ck[g] == c) ? (cs[g] >> CS_HEALTH_SHIFT) : 0u;
            if (ck[g] < c) cs[g] &= ~CS_SCORE_MASK;
        }
        clkok = cnt >= reqC;
    }
    u32 acc = 0, b = 0xFFFFFFFFu;
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const bool q = fr[g] >= m;
        b = q ? min(b, fr[g]) : b;
        acc += q ? cs[g] : 0u;
    }
    bound = b;
    fit = ok && clkok && (num <= cardnum) && ((acc >> CS_HEALTH_SHIFT) >= reqM);
    return stat + (fit ? (acc & CS_SCORE_MASK) : 0u);
}

// The pods arrive sorted by (group, memory).  Along that order the verdict and score of a node
// change only when the memory value crosses one of the node's FreeMemory values, so each lane
// keeps its last result and the bound up to which it stays valid; a 32-node slot is re-evaluated
// only when one of its lanes crossed its bound (or the group/class changed).  Within a group the
// memory values ascend, so a node that stopped fitting never fits again: its bound becomes +inf.

struct PodRegs {
    u32 m, c, num, reqM, reqC, cls, orig, gid;
    bool has_clock;
};
__device__ __forceinline__ PodRegs pod_from_smem(const uint4 *sp, u32 i) {
    const uint4 a = sp[2 * i], b = sp[2 * i + 1];  // PodS: m c num req | cls orig gid pad
    PodRegs p;
    p.m = a.x; p.c = a.y; p.num = a.z; p.reqM = a.w & 0xFFu; p.reqC = (a.w >> 8) & 0xFFu;
    p.has_clock = (a.w >> 16) & YODA_POD_HAS_CLOCK;
    p.cls = b.x; p.orig = b.y; p.gid = b.z;
    return p;
}

// Both passes are written as bodies over HOT_THREADS threads with their own shared memory and a
// named barrier, so that they can be composed (a fused output+score CTA was tried in round 1 and
// dropped: both passes need ~128 registers per thread, so they only time-share the SM).
__device__ __forceinline__ void half_sync(int bar_id) { asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(HOT_THREADS) : "memory"); }

#define SCORE_BLK 4  // chunk sizes stay multiples of this (hot_chunk in runtime.cu)
template <int NPL, bool DECIDE>
struct ScoreSmem {
    // the card contributions of the CTA's nodes for the current (group, class) run, with the group's clock tests folded in
    // (as hot_load_class does for the output pass); [half][node]: consecutive lanes read consecutive 16-byte pieces.
    // (The nodes' FreeMemory values stay in registers: the pass is bound by the shared-memory/LSU data path -- 75 % of its
    // wavefront peak with both in shared memory -- and the 16 registers still fit four CTAs per SM; score phase -5 %.)
    uint4 cs[2][HOT_WARPS * NPL * 32];
    uint4 sp[SCORE_CHUNK * 2];
    u32 m_s[SCORE_CHUNK];    // memory value of every pod of the chunk (ascending inside a run)
    u32 key_s[SCORE_CHUNK];  // run key: requirement group << 12 | class; a (group, class) run is contiguous in the sorted order
    __align__(16) u32 hi_s[HOT_WARPS][SCORE_CHUNK];
    __align__(16) u32 lo_s[HOT_WARPS][SCORE_CHUNK];
    __align__(16) u32 cnt_s[DECIDE ? HOT_WARPS : 1][DECIDE ? SCORE_CHUNK : 4];   // DECIDE only: feasible nodes of the warp's slots
    __align__(16) u32 best_s[DECIDE ? HOT_WARPS : 1][DECIDE ? SCORE_CHUNK : 4];  // DECIDE only: lowest node index holding the warp's top score
    u32 mark_s[HOT_WARPS][SCORE_CHUNK / 32];           // bit i: the warp's entry for pod i is set (it holds until the next set bit)
    u32 carry_s[HOT_WARPS][SCORE_CHUNK / 32];          // the warp's last marked pod before window k (window 0: pod 0, always marked)
};

// score reduce: lowest / highest over the feasible nodes of every pod (scheduler.go:134-145).
// The nodes' FreeMemory values live in registers and their contributions for the current run in shared memory, so
// re-evaluating a slot inside a run touches no global memory at all; only a run change fetches class-table entries (and
// clocks) from L2.
// A warp does not walk the pods one by one.  After it has evaluated its nodes for pod i, everything it knows
// stays valid up to the first later pod that either starts another (group, class) run or asks for more memory than
// the smallest validity bound among its nodes -- and because the memory values ascend inside a run, that pod is
// found by a warp-wide look-ahead over the staged values, 32 pods per step (one compare + vote), instead of a
// test per pod.  The warp's (highest, lowest) are stored only at the pods it stopped at, under mark bits; the pods
// in between inherit the entry of the last marked pod before them.
// DECIDE (per-pod results only, one GPU): the same pass also counts the feasible nodes and tracks the
// lowest-index node holding the top score, packed as (score << 32 | ~index) so that one 64-bit
// atomicMax across the node tiles yields the choice -- no output pass is needed at all.
template <int NPL, bool DECIDE>
__device__ __forceinline__ void score_body(ScoreSmem<NPL, DECIDE> &sm, int bar_id, u32 tid, u32 tile, u32 chunk, const DevSnap &s,
                                           const PodS *__restrict__ pods, u32 p_begin, u32 p_end, const u32 *__restrict__ tables,
                                           u64 *hilo, u64 *count, u64 *dkey, u32 chunk_pods) {
    const u32 lane = tid & 31, warp = tid >> 5;
    const u32 c0 = p_begin + chunk * chunk_pods;
    const u32 cn = min(chunk_pods, p_end - c0);
    for (u32 i = tid; i < cn * 2; i += HOT_THREADS) sm.sp[i] = reinterpret_cast<const uint4 *>(pods + c0)[i];
    for (u32 i = tid; i < cn; i += HOT_THREADS) {
        const PodS &ps = pods[c0 + i];
        sm.m_s[i] = ps.m;
        sm.key_s[i] = (ps.gid << FAST_GID_BITS) | ps.cls;  // classes are capped at 4096 (FAST_CLASS_CAP)
    }
    if (tid < HOT_WARPS * (SCORE_CHUNK / 32)) (&sm.mark_s[0][0])[tid] = 0u;
    const u32 base = (tile * HOT_WARPS + warp) * (NPL * 32);
    u32 node[NPL], cardnum[NPL], stat[NPL];
    bool ok[NPL], clkok[NPL];
    uint4 frv[NPL][2];  // FreeMemory of the node's eight card slots
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
        const u32 n = base + j * 32 + lane;
        const bool in = n < s.n_nodes;
        node[j] = in ? n : 0;
        cardnum[j] = s.cardnum32[node[j]];
        stat[j] = in ? s.stat32[node[j]] : 0u;
        ok[j] = in && meta_table_node(s.meta[node[j]]);
        clkok[j] = true;
        const uint4 *src = reinterpret_cast<const uint4 *>(s.n32[F_FREE] + (size_t)node[j] * G);
        frv[j][0] = src[0];
        frv[j][1] = src[1];
    }
    half_sync(bar_id);
    u32 cur_key = 0xFFFFFFFFu;
    u32 hi_l[NPL], lo_l[NPL], bound[NPL];
#pragma unroll
    for (int j = 0; j < NPL; ++j) { hi_l[j] = 0; lo_l[j] = 0xFFFFFFFFu; bound[j] = 0; }
    for (u32 i = 0; i < cn;) {  // i is warp-uniform
        const PodRegs pod = pod_from_smem(sm.sp, i);
        const u32 key = (pod.gid << FAST_GID_BITS) | pod.cls;
        const bool same = key == cur_key;
        cur_key = key;
#pragma unroll
        for (int j = 0; j < NPL; ++j) {
            const u32 slot = (warp * NPL + j) * 32 + lane;
            if (!same) {  // a new run: this slot's contributions from the class table, the group's clock tests folded in
                u32 cs[G];
                load8(tables + ((size_t)pod.cls * s.cap + node[j]) * G, cs);
                clkok[j] = true;
                if (pod.has_clock) {  // warp-uniform
                    u32 ck[G];
                    load8(s.n32[F_CLOCK] + (size_t)node[j] * G, ck);
                    u32 cnt = 0;
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        cnt += (ck[g] == pod.c) ? (cs[g] >> CS_HEALTH_SHIFT) : 0u;  // filter.go:56-58
                        if (ck[g] < pod.c) cs[g] &= ~CS_SCORE_MASK;                  // collection.go:46 / algorithm.go:48
                    }
                    clkok[j] = cnt >= pod.reqC;
                }
                sm.cs[0][slot] = make_uint4(cs[0], cs[1], cs[2], cs[3]);
                sm.cs[1][slot] = make_uint4(cs[4], cs[5], cs[6], cs[7]);
            }
            if (!same || __any_sync(0xFFFFFFFFu, pod.m > bound[j])) {
                const uint4 f0 = frv[j][0], f1 = frv[j][1], c0 = sm.cs[0][slot], c1 = sm.cs[1][slot];  // (the slot's cs entries: written by this lane only)
                const u32 fr[G] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
                const u32 cs[G] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
                u32 acc = 0, b = 0xFFFFFFFFu;
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    const bool q = fr[g] >= pod.m;  // filter.go:53 / collection.go:46 / algorithm.go:48
                    b = q ? min(b, fr[g]) : b;
                    acc += q ? cs[g] : 0u;
                }
                const bool fit = ok[j] && clkok[j] && (pod.num <= cardnum[j]) && ((acc >> CS_HEALTH_SHIFT) >= pod.reqM);
                const u32 raw = stat[j] + (fit ? (acc & CS_SCORE_MASK) : 0u);
                bound[j] = fit ? b : 0xFFFFFFFFu;  // a node that stopped fitting never fits again inside the run
                hi_l[j] = fit ? raw : 0u;
                lo_l[j] = fit ? raw : 0xFFFFFFFFu;
            }
        }
        u32 hi = hi_l[0], lo = lo_l[0], bmin = bound[0];
#pragma unroll
        for (int j = 1; j < NPL; ++j) { hi = max(hi, hi_l[j]); lo = min(lo, lo_l[j]); bmin = min(bmin, bound[j]); }
        const u32 hi_w = __reduce_max_sync(0xFFFFFFFFu, hi);
        const u32 lo_w = __reduce_min_sync(0xFFFFFFFFu, lo);
        const u32 wb = __reduce_min_sync(0xFFFFFFFFu, bmin);  // the warp's results hold for every m <= wb of this run
        u32 cnt_w = 0, best_w = 0xFFFFFFFFu;
        if (DECIDE) {
#pragma unroll
            for (int j = NPL - 1; j >= 0; --j) {
                const bool fitj = lo_l[j] != 0xFFFFFFFFu;
                cnt_w += __popc(__ballot_sync(0xFFFFFFFFu, fitj));
                const u32 top = __ballot_sync(0xFFFFFFFFu, fitj && hi_l[j] == hi_w);
                if (top) best_w = base + j * 32 + (__ffs(top) - 1);
            }
        }
        if (lane == 0) {
            sm.hi_s[warp][i] = hi_w;
            sm.lo_s[warp][i] = lo_w;
            if (DECIDE) {
                sm.cnt_s[warp][i] = cnt_w;
                sm.best_s[warp][i] = best_w;
            }
            sm.mark_s[warp][i >> 5] |= 1u << (i & 31);  // this warp's row: no other writer
        }
        // look ahead: the first later pod of another run, or beyond the bound
        u32 k = i + 1;
        for (;;) {
            const u32 idx = k + lane;
            const bool stop = idx >= cn || sm.key_s[idx] != key || sm.m_s[idx] > wb;
            const u32 bal = __ballot_sync(0xFFFFFFFFu, stop);
            if (bal) {
                k += __ffs(bal) - 1;
                break;
            }
            k += 32;
        }
        i = k;
    }
    half_sync(bar_id);
    if (tid < HOT_WARPS * (SCORE_CHUNK / 32)) {  // per (warp, window): where the warp's last mark before the window sits
        const u32 w = tid / (SCORE_CHUNK / 32), k0 = tid % (SCORE_CHUNK / 32);
        u32 k = k0, m = 0;
        while (!m && k) m = sm.mark_s[w][--k];
        sm.carry_s[w][k0] = (k << 5) + (31 - __clz(m | 1u));
    }
    half_sync(bar_id);
    for (u32 i = tid; i < cn; i += HOT_THREADS) {
        u32 hi = 0, lo = 0xFFFFFFFFu, q[HOT_WARPS];
#pragma unroll
        for (int w = 0; w < HOT_WARPS; ++w) {
            // the warp's entry for pod i sits at the last marked pod at or before i (pod 0 of a chunk is always marked)
            const u32 k = i >> 5, m = sm.mark_s[w][k] & (0xFFFFFFFFu >> (31 - (i & 31)));
            q[w] = m ? (k << 5) + (31 - __clz(m)) : sm.carry_s[w][k];
            hi = max(hi, sm.hi_s[w][q[w]]);
            lo = min(lo, sm.lo_s[w][q[w]]);
        }
        if (lo != 0xFFFFFFFFu) {  // some node of this tile is feasible
            const u32 orig = sm.sp[2 * i + 1].y;
            atomicMax(&hilo[2 * (size_t)orig], (u64)hi);
            atomicMax(&hilo[2 * (size_t)orig + 1], ~(u64)lo);
            if (DECIDE) {
                u32 cnt = 0, best = 0xFFFFFFFFu;
#pragma unroll
                for (int w = HOT_WARPS - 1; w >= 0; --w) {  // warps hold ascending node ranges: the last match wins = lowest index
                    cnt += sm.cnt_s[w][q[w]];
                    if (sm.cnt_s[w][q[w]] && sm.hi_s[w][q[w]] == hi) best = sm.best_s[w][q[w]];
                }
                atomicAdd(&count[orig], (u64)cnt);
                atomicMax(&dkey[orig], ((u64)hi << 32) | (u64)(0xFFFFFFFFu - (s.node_base + best)));
            }
        }
    }
}

struct OutSmem {
    uint4 sp[HOT_CHUNK * 2];
    uint4 sn[HOT_CHUNK];  // rescale constants per pod: lo, hi, reciprocal hi/lo words
    u32 cnt_s[HOT_WARPS][HOT_CHUNK];
    u32 best_s[HOT_WARPS][HOT_CHUNK];
    // feasible bits of the CTA's 1024 nodes for every pod of the chunk: [pod][warp] = the warp's four words.  Written out
    // after the walk as one 128-byte line per pod -- a warp's 16 bytes alone are half a DRAM sector, and half sectors
    // stored one by one cost 3.4x their share of the write stream (measured: 0.32 ms of the 6.9 ms pass for 1.4 % of its bytes)
    uint4 bits_s[HOT_CHUNK][HOT_WARPS];
};

// the rare exact rescale, kept out of line so that its 64-bit division does not bloat the hot loop
__device__ __noinline__ u32 exact_norm_u8(u32 raw, i64 lo, i64 hi) { return (u32)dev_norm_u8(dev_norm_i64((i64)raw, lo, hi)); }

// rescale constants of one pod from the (cross-shard reduced) lowest/highest: norm = ((raw-lo)*100*M) >> 57
__device__ __forceinline__ uint4 norm_consts(const u64 *__restrict__ hilo, u32 orig) {
    const u64 nlo = hilo[2 * (size_t)orig + 1];
    if (nlo == 0) return make_uint4(0, 0xFFFFFFFFu, 0, 0);  // no feasible node anywhere
    const u64 hi64 = hilo[2 * (size_t)orig], lo64 = ~nlo;
    // The 32-bit rescale below holds for lowest/highest of table nodes (scores < 2^25).  Wide nodes of a mixed snapshot,
    // or another shard on the exact kernels, can push them beyond: such a pod is marked (z = 0, w = 1) and rescaled
    // with plain 64-bit arithmetic.  y = the highest a table node can be compared with (none can reach one >= 2^32).
    if (hi64 > 0xFFFFFFFEull || hi64 - lo64 > (1ull << 25)) return make_uint4((u32)lo64, hi64 > 0xFFFFFFFEull ? 0xFFFFFFFFu : (u32)hi64, 0, 1);
    const u32 hi = (u32)hi64;
    u32 lo = (u32)lo64;
    if (hi == lo) lo = lo - 1u;  // scheduler.go:147-149 (may become -1 = 0xFFFFFFFF)
    const u32 d = hi - lo;       // 1 .. 2^25
    const u64 M = ((1ull << 57) / d) + 1ull;  // >= 2^32: z is never 0 here
    return make_uint4(lo, hi, (u32)(M >> 32), (u32)M);
}

// output pass: feasible bits, Score (int64), NormalizeScore (uint8), feasible count, node choice.
// Rows are padded to the CTA tile (stride % (HOT_WARPS*NPL*32) == 0), so no lane is out of range:
// lanes past the last node hold no SCV, score 0 and write into the padding.
template <int NPL, bool WRITE, bool RAW32>
__device__ __forceinline__ void out_body(OutSmem &sm, int bar_id, u32 tid, u32 tile, u32 chunk, const DevSnap &s,
                                         const PodS *__restrict__ pods, const u64 *__restrict__ hilo, u32 p_begin, u32 p_end,
                                         const u32 *__restrict__ tables, u64 *count, u64 *ckey, i64 *raw_out, u8 *norm_out, u32 *bits_out,
                                         u64 stride, u32 chunk_pods) {
    const u32 lane = tid & 31, warp = tid >> 5;
    const u32 c0 = p_begin + chunk * chunk_pods;
    const u32 cn = min(chunk_pods, p_end - c0);
    for (u32 i = tid; i < cn * 2; i += HOT_THREADS) sm.sp[i] = reinterpret_cast<const uint4 *>(pods + c0)[i];
    for (u32 i = tid; i < cn; i += HOT_THREADS) sm.sn[i] = norm_consts(hilo, pods[c0 + i].orig);
    const u32 base = (tile * HOT_WARPS + warp) * (NPL * 32);
    NodeRegs<NPL> r;
    hot_load_nodes(r, s, base, lane);
    half_sync(bar_id);
    u32 cur_cls = 0xFFFFFFFFu, cur_gid = 0xFFFFFFFEu, cur_m = 0, cur_lo = 0, cur_hi = 0;
    u64 raw64[NPL];                     // Score as stored (int64, high word 0)
    static_assert(NPL == 4, "the row layout below packs two pairs of adjacent nodes per lane");
    u32 bound[NPL], fitm[NPL];          // validity bound per lane; feasible lanes of slot j (warp-uniform ballots)
    u32 fitw[NPL] = {0, 0, 0, 0};       // the same bits in row order: word w = nodes base+32w .. base+32w+31
    u32 fitl[NPL], qn[NPL], rmax[NPL];  // lane mask (0 / ~0), normalized score, best feasible score of the slot
#pragma unroll
    for (int j = 0; j < NPL; ++j) { raw64[j] = 0; bound[j] = 0; fitm[j] = 0; fitl[j] = 0; qn[j] = 0; rmax[j] = 0; }
    u32 cnt = 0, bmin = 0, wmax = 0;  // bmin: min bound over the lane's slots; wmax: best feasible score of the warp's nodes
    unsigned long long *rawp = reinterpret_cast<unsigned long long *>(raw_out) + base + 2 * lane;
    u8 *normp = norm_out + base + 2 * lane;
    for (u32 i = 0; i < cn; ++i) {
        const PodRegs pod = pod_from_smem(sm.sp, i);
        const uint4 pn = sm.sn[i];  // lo, hi, mhi, mlo
        const bool same = (pod.gid == cur_gid) && (pod.cls == cur_cls) && (pod.m >= cur_m);
        const bool exact_norm = pn.z == 0 && pn.w == 1;  // lowest/highest outside the 32-bit rescale (see norm_consts)
        const bool nsame = same && (pn.x == cur_lo) && (pn.y == cur_hi) && !exact_norm;
        const bool miss_any = !same || __any_sync(0xFFFFFFFFu, pod.m > bmin);
        if (miss_any) {
            if (pod.cls != cur_cls || pod.gid != cur_gid) hot_load_class(r, s, tables, pod.cls, pod.has_clock, pod.c, pod.reqC, base, lane);
            cur_cls = pod.cls; cur_gid = pod.gid;
            cnt = 0; wmax = 0;
#pragma unroll
            for (int j = 0; j < NPL; ++j) {
                if (!same || __any_sync(0xFFFFFFFFu, pod.m > bound[j])) {
                    bool fit;
                    u32 b;
                    const u32 raw = hot_eval(r, j, pod.m, pod.num, pod.reqM, fit, b);
                    bound[j] = fit ? b : 0xFFFFFFFFu;
                    raw64[j] = raw;  // filter.go:84-86 is the identity below 2^63
                    fitm[j] = __ballot_sync(0xFFFFFFFFu, fit);
                    fitl[j] = fit ? 0xFFFFFFFFu : 0u;
                    rmax[j] = __reduce_max_sync(0xFFFFFFFFu, fit ? raw : 0u);
                    if (WRITE && nsame) {
                        const u32 x = ((u32)raw64[j] - pn.x) * 100u;
                        qn[j] = ((x * pn.z + __umulhi(x, pn.w)) >> 25) & fitl[j];
                    }
                }
                cnt += __popc(fitm[j]);
                wmax = max(wmax, rmax[j]);
            }
            bmin = bound[0];
#pragma unroll
            for (int j = 1; j < NPL; ++j) bmin = min(bmin, bound[j]);
            // node base + 64*(j>>1) + 2*lane + (j&1) is bit `lane` of ballot j: interleave pairs of ballots into row order
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                u32 word = 0;
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    u32 x = (fitm[2 * (w >> 1) + e] >> (16 * (w & 1))) & 0xFFFFu;  // lanes 16(w&1) .. +15
                    x = (x | (x << 8)) & 0x00FF00FFu;                               // spread bit i to position 2i
                    x = (x | (x << 4)) & 0x0F0F0F0Fu;
                    x = (x | (x << 2)) & 0x33333333u;
                    x = (x | (x << 1)) & 0x55555555u;
                    word |= x << e;
                }
                fitw[w] = word;
            }
        }
        if (WRITE && exact_norm) {
            i64 lo, hi;
            dev_lo_hi(hilo, pod.orig, lo, hi);
#pragma unroll  // (constant indices: a rolled loop would move qn/fitl/raw64 to local memory for every pod)
            for (int j = 0; j < NPL; ++j) qn[j] = exact_norm_u8((u32)raw64[j], lo, hi) & fitl[j];
        } else if (WRITE && !nsame) {
            // scheduler.go:152: (raw - lo) * 100 / (hi - lo), exact via the 57-bit reciprocal
#pragma unroll
            for (int j = 0; j < NPL; ++j) {
                const u32 x = ((u32)raw64[j] - pn.x) * 100u;
                qn[j] = ((x * pn.z + __umulhi(x, pn.w)) >> 25) & fitl[j];
            }
        }
        cur_m = pod.m; cur_lo = pn.x; cur_hi = pn.y;
        u32 best = 0xFFFFFFFFu;
        if (cnt && wmax == pn.y) {  // warp-uniform: some lane of this warp holds the pod's top score
            u32 mine = 0xFFFFFFFFu;
#pragma unroll
            for (int j = NPL - 1; j >= 0; --j)
                if (fitl[j] && (u32)raw64[j] == pn.y) mine = base + (j >> 1) * 64 + 2 * lane + (j & 1);
            best = __reduce_min_sync(0xFFFFFFFFu, mine);
        }
        if (WRITE) {
            const size_t row = (size_t)pod.orig * stride;
            // streaming stores (written once, never re-read by this pass); every instruction covers whole sectors:
            // 2 x 512 B of scores and 2 x 64 B of normalized scores per warp
            if (RAW32) {  // YODA_EVAL_RAW32: the same values without their zero high words, 2 x 256 B per warp
                u32 *rawp32 = reinterpret_cast<u32 *>(raw_out) + base + 2 * lane;
                __stcs(reinterpret_cast<uint2 *>(rawp32 + row), make_uint2((u32)raw64[0], (u32)raw64[1]));
                __stcs(reinterpret_cast<uint2 *>(rawp32 + row + 64), make_uint2((u32)raw64[2], (u32)raw64[3]));
            } else {
                __stcs(reinterpret_cast<ulonglong2 *>(rawp + row), make_ulonglong2(raw64[0], raw64[1]));
                __stcs(reinterpret_cast<ulonglong2 *>(rawp + row + 64), make_ulonglong2(raw64[2], raw64[3]));
            }
            __stcs(reinterpret_cast<unsigned short *>(normp + row), (unsigned short)(qn[0] | (qn[1] << 8)));
            __stcs(reinterpret_cast<unsigned short *>(normp + row + 64), (unsigned short)(qn[2] | (qn[3] << 8)));
        }
        if (lane == 0) {
            sm.cnt_s[warp][i] = cnt;
            sm.best_s[warp][i] = best;
            if (WRITE) sm.bits_s[i][warp] = make_uint4(fitw[0], fitw[1], fitw[2], fitw[3]);
        }
    }
    half_sync(bar_id);
    if (WRITE) {  // the chunk's feasible-bit rows: warp w takes pods w, w + 8, ...; lane = word of the CTA's 32-word piece
        u32 *piece = bits_out + (size_t)tile * (HOT_WARPS * NPL) + lane;
        for (u32 i = warp; i < cn; i += HOT_WARPS) {
            const u32 orig = sm.sp[2 * i + 1].y;
            __stcs(piece + (size_t)orig * (stride >> 5), reinterpret_cast<const u32 *>(&sm.bits_s[i][0])[lane]);
        }
    }
    for (u32 i = tid; i < cn; i += HOT_THREADS) {
        u32 cnt2 = 0, best = 0xFFFFFFFFu;
#pragma unroll
        for (int w = 0; w < HOT_WARPS; ++w) {
            cnt2 += sm.cnt_s[w][i];
            best = min(best, sm.best_s[w][i]);
        }
        const u32 orig = sm.sp[2 * i + 1].y;
        if (cnt2) atomicAdd(&count[orig], (u64)cnt2);
        if (best != 0xFFFFFFFFu) atomicMax(&ckey[orig], ckey_of(s.node_base + best));
    }
}

#ifndef SCORE_MINB
#define SCORE_MINB 4  // resident CTAs per SM the score pass is compiled for (62 registers, ~43 KB of shared memory)
#endif
template <int NPL, bool DECIDE>
__global__ void __launch_bounds__(HOT_THREADS, SCORE_MINB) k_fast_score(DevSnap s, const PodS *__restrict__ pods, u32 p_begin, u32 p_end,
                                                            const u32 *__restrict__ tables, u64 *hilo, u64 *count, u64 *dkey,
                                                            u32 chunk_pods,  // pods per CTA: a multiple of SCORE_BLK, <= SCORE_CHUNK
                                                            const u32 *gate) {  // non-null: run only while *gate == 0 (see gate_closed)
    extern __shared__ __align__(16) unsigned char score_smem_raw[];  // > 48 KB: dynamic (cudaFuncSetAttribute in yoda_create)
    ScoreSmem<NPL, DECIDE> &sm = *reinterpret_cast<ScoreSmem<NPL, DECIDE> *>(score_smem_raw);
    if (gate_closed(gate, gate ? GATE_IF_ZERO : GATE_NONE)) return;
    score_body<NPL, DECIDE>(sm, 1, threadIdx.x, blockIdx.x, blockIdx.y, s, pods, p_begin, p_end, tables, hilo, count, dkey, chunk_pods);
}

template <int NPL, bool WRITE, bool RAW32>
__global__ void __launch_bounds__(HOT_THREADS, 2) k_fast_out(DevSnap s, const PodS *__restrict__ pods, const u64 *__restrict__ hilo,
                                                             u32 p_begin, u32 p_end, const u32 *__restrict__ tables, u64 *count, u64 *ckey,
                                                             i64 *raw_out, u8 *norm_out, u32 *bits_out, u64 stride, u32 chunk_pods,
                                                             const u32 *gate) {  // non-null: run only while *gate == 0
    __shared__ OutSmem sm;
    if (gate_closed(gate, gate ? GATE_IF_ZERO : GATE_NONE)) return;
    out_body<NPL, WRITE, RAW32>(sm, 1, threadIdx.x, blockIdx.x, blockIdx.y, s, pods, hilo, p_begin, p_end, tables, count, ckey, raw_out, norm_out,
                         bits_out, stride, chunk_pods);
}
