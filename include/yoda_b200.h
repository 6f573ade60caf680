/*
 * yoda_b200.h -- C ABI of the B200-native Yoda Filter/Score hot path (libyoda_b200.so).
 *
 * This is the drop-in boundary: plain C types, opaque handle, no C++/torch types.  It is what
 * a cgo shim inside the reference's pkg/yoda binds.  Every entry point
 * cites the reference interface it replaces; paths are relative to the reference checkout
 * (github.com/NJUPT-ISL/Yoda-Scheduler).
 *
 * Conventions
 *   - every function returns YODA_OK (0) or a negative YODA_E_* code and never aborts or
 *     throws; yoda_last_error() gives the thread-local message of the last failure.
 *   - the library owns all device memory and its pinned staging memory; caller buffers are
 *     read/written only during the call (cgo rule: no caller pointer is retained).
 *   - one mutex per context: upsert/commit/eval calls on one context are serialised.
 *   - there is NO CPU fallback: without a usable CUDA device yoda_create fails.
 *   - node indices are dense, caller-assigned (node name <-> index lives on the host side).
 */
#ifndef YODA_B200_H
#define YODA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YODA_ABI_VERSION 1
#define YODA_MAX_CARDS 8 /* cards per node held in the device snapshot (len(Status.CardList) <= 8) */

/* error codes */
#define YODA_OK 0
#define YODA_E_ARG (-1)    /* bad argument */
#define YODA_E_CUDA (-2)   /* CUDA runtime error (message has the CUDA string) */
#define YODA_E_RANGE (-3)  /* index / size outside the configured capacity */
#define YODA_E_NOMEM (-4)  /* allocation failed */
#define YODA_E_STATE (-5)  /* call order violated (e.g. eval with uncommitted flushed upserts) */
#define YODA_E_DIV0 (-6)   /* TotalMemorySum == 0: the reference would panic (algorithm.go:71,86) */
#define YODA_E_NCCL (-7)   /* NCCL unavailable or failed */

typedef struct yoda_ctx yoda_ctx;

typedef struct {
    int32_t device;       /* CUDA device ordinal */
    uint32_t max_nodes;   /* capacity of this context's (shard of the) snapshot */
    uint32_t node_base;   /* global index of local node 0 (node-sharded configs, else 0) */
    uint32_t shard_rank;  /* this shard's rank, 0 when unsharded */
    uint32_t shard_count; /* number of node shards, 1 when unsharded */
    uint32_t flags;       /* reserved, 0 */
} yoda_config;

/* One node's SCV object: the fields of scv.Scv / scv.Card the hot path reads (SURVEY.md 8(a) T1;
 * call sites filter.go:13-58, collection.go:46-77, algorithm.go:58-86).  Arrays are indexed by
 * position in Status.CardList.  healthy_mask bit g = (CardList[g].Health == "Healthy"). */
typedef struct {
    uint64_t card_number;      /* Status.CardNumber (independent of n_cards) */
    uint64_t free_memory_sum;  /* Status.FreeMemorySum */
    uint64_t total_memory_sum; /* Status.TotalMemorySum, must be > 0 */
    uint32_t n_cards;          /* len(Status.CardList), <= YODA_MAX_CARDS */
    uint32_t healthy_mask;
    uint64_t free_memory[YODA_MAX_CARDS];
    uint64_t total_memory[YODA_MAX_CARDS];
    uint64_t clock[YODA_MAX_CARDS];
    uint64_t core[YODA_MAX_CARDS];
    uint64_t bandwidth[YODA_MAX_CARDS];
    uint64_t power[YODA_MAX_CARDS];
} yoda_node_record;

/* One pending pod's requirement vector: its scv/number, scv/memory, scv/clock labels after
 * strconv.Atoi-with-error-to-zero (filter.go:60-74; use yoda_parse_label).  32 bytes. */
#define YODA_POD_HAS_NUMBER 1u
#define YODA_POD_HAS_MEMORY 2u
#define YODA_POD_HAS_CLOCK 4u
typedef struct {
    uint32_t flags; /* YODA_POD_HAS_*: label present in pod.GetLabels() */
    uint32_t reserved;
    uint64_t number, memory, clock;
} yoda_pod_req;

/* Per-pod outcome of one scheduling cycle.  72 bytes. */
typedef struct {
    /* collection.MaxValue in its declaration order (collection.go:14-21) */
    uint64_t max_bandwidth, max_clock, max_core, max_free_memory, max_power, max_total_memory;
    int64_t lowest, highest; /* as used by the rescale at scheduler.go:152 (after lowest--) */
    uint32_t feasible;       /* number of nodes that passed Filter (all shards) */
    int32_t chosen;          /* global index of the lowest-index node with the top score, -1 if none */
} yoda_pod_result;

/* eval flags */
#define YODA_EVAL_DECIDE 0u       /* per-pod results only */
#define YODA_EVAL_FULL 1u         /* also materialise feasible bits, raw and normalized scores in HBM */
#define YODA_EVAL_FORCE_DIRECT 2u /* use the exact-u64 direct kernels even when the narrow path applies */
/* With YODA_EVAL_FULL: store the Score matrix as uint32 instead of int64.  Every Score of a narrow snapshot is
 * below 2^32 (it is what the fast path computes in), so this only drops the 32 zero bits of each int64 and with
 * them almost half of the bytes the output pass writes (9.125 -> 5.125 per pod x node).  A request, not a
 * guarantee: a batch that has to take the exact-u64 kernels still gets int64 rows; yoda_result_raw_bytes tells
 * which, and yoda_fetch_row(YODA_ROW_RAW) delivers int64 either way. */
#define YODA_EVAL_RAW32 4u

/* rows of the device-resident result matrices (YODA_EVAL_FULL) */
#define YODA_ROW_FEASIBLE_BITS 0 /* uint32 words, bit (n & 31) of word n >> 5 = node n passed Filter */
#define YODA_ROW_RAW 1           /* int64 per node: what Yoda.Score returns (0 for nodes without SCV) */
#define YODA_ROW_NORM 2          /* uint8 per node: NormalizeScore output, 0 for infeasible nodes */

/* Device time of the phases of the last eval (CUDA events on the context's stream), ms. */
typedef struct {
    float total_ms;     /* first kernel -> last kernel, copies excluded */
    float prep_ms;      /* pod canonicalisation, ordering, class tables */
    float max_ms;       /* collection.CollectMaxValues */
    float score_ms;     /* Filter+Score reduce (lowest/highest) */
    float write_ms;     /* Filter+Score+NormalizeScore output pass (node choice) */
    float h2d_ms, d2h_ms;
    uint32_t kernels;   /* kernels launched by the last eval */
    uint32_t path;      /* 0 = direct exact-u64, 1 = narrow fast path */
    uint32_t n_classes; /* distinct score-divisor tuples in the batch (fast path) */
    uint32_t n_groups;  /* distinct (clock, number, has-memory) requirement groups (fast path) */
} yoda_timing;

/* ---- lifecycle: replaces yoda.New's cache construction (pkg/yoda/scheduler.go:46-74) ---- */
int yoda_create(const yoda_config *cfg, yoda_ctx **out);
void yoda_destroy(yoda_ctx *ctx);
int yoda_abi_version(void);
const char *yoda_last_error(void);

/* ---- snapshot maintenance: replaces the controller-runtime informer cache reads
 *      y.cache.Get / y.cache.List (scheduler.go:80,98,118).  Called from Scv add/update/delete
 *      and Pod add/delete event handlers.  Upserts are staged in pinned memory and become
 *      visible to evaluations at the next yoda_commit. ---- */
int yoda_node_upsert(yoda_ctx *ctx, uint32_t idx, const yoda_node_record *rec);
int yoda_node_upsert_many(yoda_ctx *ctx, uint32_t n, const uint32_t *idx, const yoda_node_record *recs);
/* sum of StrToUint64(scv/memory) over the pods running on the node (algorithm.go:75-80) */
int yoda_node_set_alloc(yoda_ctx *ctx, uint32_t idx, uint64_t allocate_memory_sum);
int yoda_node_set_alloc_many(yoda_ctx *ctx, uint32_t n, const uint32_t *idx, const uint64_t *allocate_memory_sum);
/* Scv deleted: cache.Get fails -> node Unschedulable (scheduler.go:81-84) */
int yoda_node_remove(yoda_ctx *ctx, uint32_t idx);
int yoda_commit(yoda_ctx *ctx, uint64_t *epoch);
int yoda_node_count(yoda_ctx *ctx, uint32_t *n_nodes); /* highest committed index + 1 */

/* ---- evaluation: replaces, for a whole batch of pending pods on one frozen snapshot,
 *      Yoda.Filter (scheduler.go:76-93) for every node, Yoda.PostFilter/CollectMaxValues
 *      (scheduler.go:95-103, collection.go:30-57), Yoda.Score (scheduler.go:109-130) and
 *      Yoda.NormalizeScore (scheduler.go:132-157), plus the lowest-index argmax. ---- */
/* out may be NULL: the batch is evaluated and its matrices kept, but no per-pod results are copied back.  For node shards
 * whose host does not consume them -- after the exchanges every shard holds the same per-pod results, one copy is enough. */
int yoda_eval_batch(yoda_ctx *ctx, const yoda_pod_req *pods, uint32_t n_pods, uint32_t flags, yoda_pod_result *out);
/* same, with `d_pods` (n_pods * sizeof(yoda_pod_req)) and `d_out` (n_pods * sizeof(yoda_pod_result),
 * may be NULL) already in device memory.  Everything is queued on the context's stream; the call returns once the
 * verdicts the kernels left on the device (which path ran, capacity overflows) have been read back -- i.e. after the
 * batch's kernels, but without copying any results -- and has by then redone the batch if they asked for it. */
int yoda_eval_batch_device(yoda_ctx *ctx, const void *d_pods, uint32_t n_pods, uint32_t flags, void *d_out);
/* copy one row of the last YODA_EVAL_FULL result to the host: the O(1)-per-node memo behind
 * later Filter/Score/NormalizeScore calls of the same cycle */
int yoda_fetch_row(yoda_ctx *ctx, uint32_t pod, int kind, void *dst, size_t bytes);
/* the three rows of one pod in one call (three DMA copies through pinned staging, one synchronisation): what the
 * plugin memoises at the first Filter of a cycle.  bits: (n_nodes+31)/32 words, raw: n_nodes int64, norm: n_nodes bytes */
int yoda_fetch_rows(yoda_ctx *ctx, uint32_t pod, uint32_t n_nodes, uint32_t *bits, int64_t *raw, uint8_t *norm);
int yoda_sync(yoda_ctx *ctx);
int yoda_last_timing(yoda_ctx *ctx, yoda_timing *t);
/* device pointers of the last YODA_EVAL_FULL matrices (for zero-copy consumers); row strides in elements */
int yoda_result_matrices(yoda_ctx *ctx, void **d_bits, void **d_raw, void **d_norm, uint64_t *stride_nodes);
/* element size of the raw matrix of the last YODA_EVAL_FULL result: 8 (int64), or 4 (uint32, YODA_EVAL_RAW32 honoured) */
int yoda_result_raw_bytes(yoda_ctx *ctx, uint32_t *bytes);

/* Introspection for tests and profiling (no counterpart in the reference).  After an evaluation that took the
 * narrow table path: position k of the (group, memory)-sorted walk of the hot kernels -> index of that pod in the
 * caller's batch, its requirement group and its class (any of the three arrays may be NULL; n = pods of that
 * batch), and the pods per CTA the score pass and the output pass ran with.  The parity tests use it to aim at
 * the pods on chunk / group / class boundaries.  YODA_E_STATE if the last evaluation took the exact kernels. */
int yoda_debug_pod_order(yoda_ctx *ctx, uint32_t *orig, uint32_t *group, uint32_t *cls, uint32_t n, uint32_t *score_chunk,
                         uint32_t *out_chunk);

/* Yoda.NormalizeScore on an arbitrary score list, in place, on the device (scheduler.go:132-157):
 * used when the framework's list is not Yoda's own feasible set (other plugins filtered too). */
int yoda_normalize_scores(yoda_ctx *ctx, int64_t *scores, uint32_t n);

/* Sequential-commit batch: what a run of single-pod cycles does.  Pods are taken in `order` (NULL =
 * 0..n-1, e.g. the output of yoda_queue_order); after each choice the chosen node's allocateMemorySum
 * grows by the pod's StrToUint64(scv/memory) -- kube-scheduler's assume step, which the next cycle's
 * CalculateAllocateScore sees through NodeInfo.Pods (algorithm.go:74-87) -- before the next pod is
 * evaluated.  out[p] (indexed like `pods`) is the pod's cycle on the state it met: maxima, lowest,
 * highest, feasible count, chosen node.  The committed snapshot is not modified: the caller records
 * real bindings with yoda_node_set_alloc.  Needs a narrow snapshot and an unsharded context. */
int yoda_eval_sequential(yoda_ctx *ctx, const yoda_pod_req *pods, uint32_t n_pods, const uint32_t *order, yoda_pod_result *out);

/* QueueSort for a pending batch: sort.Less / sort.GetPodPriority (pkg/yoda/sort/sort.go:8-18, wired at
 * scheduler.go:105-107) order pods by the scv/priority label, higher first.  Given the priorities
 * (host parses the label: strconv.Atoi, value kept even on a range error, 0 without the label), writes
 * the batch order: order[k] = index of the pod consumed k-th.  Pods of equal priority keep their
 * arrival order (Less is false both ways for them). */
int yoda_queue_order(yoda_ctx *ctx, const int64_t *priority, uint32_t n, uint32_t *order);

/* ---- framework tail: what kube-scheduler v1.20 does around the plugin for each pod, applied on the
 *      device to the matrices of the last YODA_EVAL_FULL batch -- for the stand-alone batch mode; behind the real
 *      framework these steps stay the framework's.  Replaces, per pod: numFeasibleNodesToFind + the round-robin
 *      node walk of findNodesThatPassFilters (percentageOfNodesToScore, deploy/yoda-scheduler.yaml:18),
 *      Yoda.NormalizeScore over the nodes found (scheduler.go:132-157), the plugin weight (deploy/yoda-scheduler.yaml:30)
 *      plus the other plugins' scores, and selectHost's reservoir tie-break with a seeded generator in place of
 *      math/rand (cmd/scheduler/main.go:13 seeds it from the clock).  Nodes are taken strictly in visiting order,
 *      one of the outcomes upstream's 16 filter goroutines can produce. ---- */
#define YODA_TAIL_CHAIN 1u /* pods are consecutive cycles: each starts its node walk where the previous one stopped */
typedef struct {
    uint32_t weight;       /* plugin weight; 300 in the shipped profile */
    int32_t percentage;    /* percentageOfNodesToScore; <= 0 = adaptive (50 - nodes/125, at least 5) */
    uint64_t seed;         /* tie-break generator: rand.Intn(k) := SplitMix64(seed, pod, k) % k */
    uint32_t start_index;  /* nextStartNodeIndex of the first (YODA_TAIL_CHAIN) or of every pod */
    uint32_t flags;        /* YODA_TAIL_* */
    const int64_t *other_scores; /* host, one per node: what the other plugins add to a node's total; NULL = 0 */
} yoda_tail_config;
typedef struct {
    int32_t chosen;          /* global node index, -1 if no node passed Filter */
    uint32_t sampled;        /* feasible nodes found (<= numFeasibleNodesToFind) */
    uint32_t processed;      /* nodes visited; the next cycle starts at (start + processed) % nodes */
    uint32_t start;          /* where this pod's walk started */
    int64_t lowest, highest; /* of the nodes found, as NormalizeScore used them (0, 0 when scoring was skipped: < 2 nodes) */
    int64_t top;             /* winning total = normalized * weight + other */
    uint32_t ties, reserved; /* nodes holding the winning total */
} yoda_tail_result;
uint32_t yoda_num_feasible_nodes_to_find(uint32_t num_all_nodes, int32_t percentage);
/* out: one entry per pod of the last YODA_EVAL_FULL batch; next_start (may be NULL): nextStartNodeIndex afterwards.
 * Unsharded contexts only. */
int yoda_framework_tail(yoda_ctx *ctx, const yoda_tail_config *cfg, yoda_tail_result *out, uint32_t *next_start);

/* No counterpart in the reference.  The same rescale for callers that may not allocate: no
 * context, no cudaMalloc/cudaFree, no host allocation -- the only scratch is a fixed 16 KiB array
 * inside the library image (resident on every GPU that loads the library, until exit), which
 * limits the list to YODA_NOALLOC_MAX_SCORES.  d_scores_in / d_scores_out are device pointers on
 * the current device (they may be equal); the work is queued on cuda_stream (cudaStream_t, NULL =
 * default stream) and is not waited for.  The first call on each device looks up the scratch
 * address and makes the CUDA runtime load the module that holds it; make that call at start-up,
 * before device memory runs short and before any stream capture begins.  Calls on one device share the
 * scratch: calls whose kernels overlap on the GPU race, so serialise them.  Longer lists are
 * rejected with YODA_E_RANGE. */
#define YODA_NOALLOC_MAX_SCORES (1u << 20)
int yoda_normalize_scores_noalloc(const int64_t *d_scores_in, int64_t *d_scores_out, uint32_t n, void *cuda_stream);

/* Host only: an Scv custom resource as JSON (the whole object, or its status alone) -> yoda_node_record; the ingest step
 * of SURVEY.md 8(f)4.  Replaces the decoding the controller-runtime client does before y.cache.Get hands out a *scv.Scv
 * (scheduler.go:79-80).  Fields read: status.{cardNumber, freeMemorySum, totalMemorySum, cardList[]} and per card {health,
 * freeMemory, totalMemory, clock, core, bandwidth, power}; the Go types live in the un-vendored SCV module, so key names
 * are matched without regard to case and unknown keys are skipped.  healthy bit g = (health == "Healthy"), as
 * filter.go:53,57 compare.  name (may be NULL): receives metadata.name.  YODA_E_ARG: malformed JSON or no status fields;
 * YODA_E_RANGE: more than YODA_MAX_CARDS cards (rec->n_cards holds the count). */
int yoda_record_from_scv_json(const char *json, size_t len, yoda_node_record *rec, char *name, size_t name_cap);
/* filter.StrToUint64 / strToUint (filter.go:60-74): strconv.Atoi, any error -> 0 */
uint64_t yoda_parse_label(const char *s, size_t len);
/* filter.Uint64ToInt64 (filter.go:84-86): values >= 2^63 -> 0 */
int64_t yoda_uint64_to_int64(uint64_t v);

/* ---- node-sharded evaluation (BASELINE config 4): each shard holds nodes
 *      [node_base, node_base + n) and sees every pod.  Either let the library reduce the
 *      per-pod partials over NCCL (yoda_comm_*), or drive the phases yourself and reduce the
 *      exchange buffers with any collective (yoda_phase_*). ---- */
#define YODA_NCCL_ID_BYTES 128
int yoda_comm_unique_id(void *id128);                                                    /* rank 0 */
int yoda_comm_init(yoda_ctx *ctx, const void *id128, uint32_t rank, uint32_t n_ranks); /* all ranks */
/* With a communicator the library runs the three per-pod exchanges of a batch itself: over NCCL by default, or -- environment
 * YODA_B200_XCHG=peer -- with its own kernels over CUDA-IPC mapped peer memory (one process per GPU on one box; falls back
 * to NCCL on every shard if any shard cannot map its peers).  yoda_destroy of such a context is collective. */

#define YODA_XCHG_MAX 0
#define YODA_XCHG_SUM 1
typedef struct {
    void *d_ptr;    /* device pointer, uint64 elements */
    uint64_t count; /* elements */
    int32_t op;     /* YODA_XCHG_MAX / YODA_XCHG_SUM over shards, element-wise, unsigned */
    int32_t reserved;
} yoda_xchg;
/* phases: 0 = upload pods + local maxima; 1 = local lowest/highest/count; 2 = local output pass
 * and node choice; 3 = finish (write results).  After phases 0, 1 and 2 the buffers returned by
 * yoda_phase_exchange must be reduced across shards before the next phase runs
 * (phase 0: the six maxima; phase 1: highest / ~lowest; phase 2: choice keys (MAX) and feasible counts (SUM)). */
int yoda_phase_begin(yoda_ctx *ctx, const yoda_pod_req *pods, uint32_t n_pods, uint32_t flags);
int yoda_phase_run(yoda_ctx *ctx, int phase);
int yoda_phase_exchange(yoda_ctx *ctx, int phase, yoda_xchg *bufs, int *n_bufs /* in: capacity, out: used */);
int yoda_phase_finish(yoda_ctx *ctx, yoda_pod_result *out);

/* run subsequent work of this context on the caller's CUDA stream (cudaStream_t), NULL = own stream */
int yoda_set_stream(yoda_ctx *ctx, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* YODA_B200_H */
