// xchg_kernels.cuh -- the per-pod exchanges of a node-sharded batch over peer memory (NVLink / NVSwitch), without NCCL.
//
// Every shard owns an ARENA in its HBM that the other shards of the box map through CUDA IPC.  An exchange of one
// partial buffer (element-wise MAX or SUM over the shards, uint64) is two small kernels on the shard's stream:
//
//   k_x_publish   copies the shard's partial values into its arena; the block that finishes last raises this shard's
//                 flag in EVERY shard's arena (a plain store over NVLink after a system-scope fence);
//   k_x_reduce    waits until every shard's flag for this exchange is up in its own arena, then reads the element from
//                 every arena (peer loads) and writes the reduced value into the shard's private buffer.
//
// One-shot: every shard reduces everything itself -- the payloads are <= 4.8 MB and the cost is latency, not bytes.
// No shard ever writes another shard's data, only its flags.  The arenas are double-buffered by batch parity: a shard
// can be at most one exchange ahead of the slowest one (it cannot pass an exchange before everyone has published for
// it), so the values a slow shard is still reading are never the ones a fast shard overwrites -- no trailing barrier.
// Flags carry the batch sequence number, which only grows, so nothing is ever cleared.  A shard that waits longer than
// ~2 s gives up and raises an error word instead of hanging the GPU.
//
// Peer memory is touched with ordinary loads and stores only.
#pragma once
#include "common.cuh"

#define X_MAX_RANKS 8
#define X_PHASES 3
#define X_WORDS_PER_POD 10  // max6 (6) + hilo (2) + ckey (1) + count (1)

struct XPeers {
    u64 *arena[X_MAX_RANKS];  // every shard's arena as seen from this shard (own included)
};

// layout of an arena holding `cap` pods per parity:
//   [parity 0 values: cap * 10 + 8 batch-wide words][parity 1 values: the same][flags: 2 * X_PHASES * X_MAX_RANKS]
#define X_EXTRA_WORDS 8  // per parity, after the per-pod values; word 0 = "a shard hit a capacity limit" (MAX)
__host__ __device__ __forceinline__ size_t x_parity_words(u32 cap) { return (size_t)cap * X_WORDS_PER_POD + X_EXTRA_WORDS; }
__host__ __device__ __forceinline__ size_t x_values_off(u32 cap, u32 parity) { return (size_t)parity * x_parity_words(cap); }
__host__ __device__ __forceinline__ size_t x_flags_off(u32 cap, u32 parity, u32 phase) {
    return 2 * x_parity_words(cap) + ((size_t)parity * X_PHASES + phase) * X_MAX_RANKS;
}
__host__ __device__ __forceinline__ size_t x_arena_words(u32 cap) { return 2 * x_parity_words(cap) + 2 * X_PHASES * X_MAX_RANKS; }

// src -> own arena.  signal != 0: this is the last buffer of the exchange; the block finishing last tells every shard.
__global__ void k_x_publish(const u64 *__restrict__ src, u64 *dst, u64 n, u32 *done, XPeers peers, u64 flags_off, u32 my_rank, u32 n_ranks,
                            u64 seq, int signal) {
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
    if (!signal) return;
    __threadfence_system();  // this thread's values are visible to the peers before the count below says so
    __syncthreads();
    if (threadIdx.x == 0) {
        const u32 prev = atomicAdd(done, 1u);
        if (prev == gridDim.x - 1) {  // every block of this launch (and, by stream order, of the earlier buffers) is done
            *done = 0;
            __threadfence_system();
            for (u32 r = 0; r < n_ranks; ++r) *reinterpret_cast<volatile u64 *>(peers.arena[r] + flags_off + my_rank) = seq;
        }
    }
}

// op: 0 = MAX, 1 = SUM.  my_flags: this shard's own flag row for the exchange.  err: set when a peer never arrived.
__global__ void k_x_reduce(XPeers peers, u64 values_off, u64 *out, u64 n, int op, const u64 *my_flags, u32 n_ranks, u64 seq, u32 *err) {
    __shared__ int s_bad;
    if (threadIdx.x == 0) {
        int bad = 0;
        const long long t0 = clock64();
        for (u32 r = 0; r < n_ranks && !bad; ++r) {
            while (*reinterpret_cast<const volatile u64 *>(my_flags + r) < seq) {
                if (clock64() - t0 > 4000000000ll) {  // ~2 s at 1.9 GHz: a peer is gone; do not hang the GPU
                    bad = 1;
                    break;
                }
                __nanosleep(200);
            }
        }
        __threadfence_system();  // the peers' values were published before their flags
        if (bad) atomicExch(err, 1u);
        s_bad = bad;
    }
    __syncthreads();
    if (s_bad) return;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        u64 acc = 0;  // identity of both MAX (unsigned) and SUM
        for (u32 r = 0; r < n_ranks; ++r) {
            const u64 v = __ldcv(peers.arena[r] + values_off + i);  // peer memory: never from a stale cache line
            acc = op ? acc + v : max(acc, v);
        }
        out[i] = acc;
    }
}
