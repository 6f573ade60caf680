// common.cuh -- shared types of the device snapshot and the per-pod records.
//
// Device snapshot ("wide planes", the source of truth in HBM): one plane per scv.Card field,
// node-major with the YODA_MAX_CARDS card slots of a node contiguous, uint64 so that every
// value the reference can hold (Go uint/uint64) is representable.  Narrow uint32 planes for
// the fast path are derived from these, node by node, right after an update is scattered (k_prepare_updates).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/yoda_b200.h"

typedef unsigned long long u64;
typedef long long i64;
typedef uint32_t u32;
typedef int32_t i32;
typedef uint8_t u8;

#define G YODA_MAX_CARDS

// node meta word: bits 0-7 healthy mask, bits 8-11 n_cards, bit 16 has_scv
#define META_HAS_SCV 0x10000u
#define META_WIDE 0x20000u  // derived: the node (with an SCV) has a value outside the narrow range; counted in the snapshot-wide verdict
// bits 24-31 (derived at commit): cards that no other card of the node dominates in all six fields -- the only ones
// CollectMaxValues can ever take a maximum from
// a node the narrow table path evaluates: it has an SCV and every value of it fits the narrow planes.  Wide nodes of a
// mixed snapshot are left to the exact kernels (k_direct_pods over DevSnap-external wide list), in the same batch.
__host__ __device__ __forceinline__ bool meta_table_node(u32 m) { return (m & (META_HAS_SCV | META_WIDE)) == META_HAS_SCV; }
__host__ __device__ __forceinline__ u32 meta_keep(u32 m) { return m >> 24; }
__host__ __device__ __forceinline__ u32 meta_healthy(u32 m) { return m & 0xFFu; }
__host__ __device__ __forceinline__ u32 meta_ncards(u32 m) { return (m >> 8) & 0xFu; }

enum Field { F_FREE = 0, F_TOTAL = 1, F_CLOCK = 2, F_CORE = 3, F_BW = 4, F_POWER = 5, NFIELD = 6 };

struct DevSnap {
    u64 *f[NFIELD];     // [cap][G] card planes
    u64 *card_number;   // [cap] Status.CardNumber
    u64 *free_sum;      // [cap] Status.FreeMemorySum
    u64 *total_sum;     // [cap] Status.TotalMemorySum
    u64 *alloc_sum;     // [cap] sum of running pods' scv/memory (algorithm.go:75-80)
    u64 *stat;          // [cap] allocate score + actual score (algorithm.go:70-87), derived
    u32 *meta;          // [cap]
    // narrow planes, derived (valid when the snapshot is narrow-eligible)
    u32 *n32[NFIELD];   // [cap][G] uint32 copies of the card planes
    u32 *cardnum32;     // [cap]
    u32 *stat32;        // [cap]
    u32 *tk32;          // [cap][G] tk32[n][k] = (k+1)-th largest FreeMemory among the healthy cards (0 if fewer)
    // nodes ordered by the Clock of their first card (nodes without an SCV last): neighbouring lanes then agree on
    // an scv/clock test, so the per-group work of the maxima pass is warp-uniform.  Rebuilt at commit when a key changed.
    u32 *order;         // [cap] permutation of 0..n_nodes-1
    u32 *order_key;     // [cap] the key each node was last ordered by
    u32 n_nodes;        // committed node count (highest index + 1)
    u32 cap;
    u32 node_base;      // global index of local node 0
};

// canonical pod record built by k_pod_prep from yoda_pod_req
//   num  = scv/number present ? value : 1                       (filter.go:11-16)
//   mem  = scv/memory present ? value : 0                       (filter.go:18-33)
//   clk  = scv/clock  present ? value : 0                       (filter.go:35-50)
struct PodC {
    u64 num, mem, clk;
    u32 flags;
    u32 pad;
};

// update record staged in pinned host memory and scattered into the planes at commit
#define UPD_REC 1u     // rec valid: upsert
#define UPD_ALLOC 2u   // alloc valid
#define UPD_REMOVE 4u  // Scv deleted
struct UpdRec {
    u32 idx;
    u32 op;
    u64 alloc;
    yoda_node_record rec;
};

// u64 order-preserving encodings used in the per-pod partial buffers
//   hilo[2p]   = highest (>= 0)                    identity 0
//   hilo[2p+1] = ~lowest over feasible nodes       identity 0  ("no feasible node yet")
//   ckey[p]    = 2^32 - global node index          identity 0  (MAX picks the lowest index)
__host__ __device__ __forceinline__ u64 ckey_of(u32 gidx) { return (u64)(0xFFFFFFFFu - gidx) + 1ull; }
__host__ __device__ __forceinline__ i32 ckey_to_idx(u64 k) { return k == 0 ? -1 : (i32)(0xFFFFFFFFu - (u32)(k - 1ull)); }
