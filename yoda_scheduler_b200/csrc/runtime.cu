// runtime.cu -- host runtime behind the C ABI of include/yoda_b200.h:
// device snapshot, pinned update staging, eval workspace, phase driver, NCCL exchange.
#include <dlfcn.h>
#include <nccl.h>  // types only; the library is dlopen'ed on demand (yoda_comm_*)

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include <cub/device/device_radix_sort.cuh>  // library utility: orders the (<= few 100k) pod keys, off the hot path

#include "common.cuh"
#include "direct_kernels.cuh"
#include "fast_kernels.cuh"
#include "seq_kernels.cuh"
#include "tail_kernels.cuh"
#include "xchg_kernels.cuh"

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
thread_local char yoda_g_err[512] = "";  // also written by noalloc.cu
#define g_err yoda_g_err
static int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}
#define CU(x)                                                                                              \
    do {                                                                                                   \
        cudaError_t e_ = (x);                                                                              \
        if (e_ != cudaSuccess) return fail(YODA_E_CUDA, "%s failed: %s (%s:%d)", #x, cudaGetErrorString(e_), \
                                           __FILE__, __LINE__);                                            \
    } while (0)
#define TRY(x)              \
    do {                    \
        int rc_ = (x);      \
        if (rc_) return rc_; \
    } while (0)

// ------------------------------------------------------------------------------------------
// NCCL through dlopen (the library must load and run on one GPU without NCCL present)
// ------------------------------------------------------------------------------------------
struct NcclApi {
    void *h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mu;
static int nccl_load() {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.h) return YODA_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(YODA_E_NCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
#define SYM(n)                                                         \
    *(void **)(&g_nccl.n) = dlsym(h, "nccl" #n);                       \
    if (!g_nccl.n) return fail(YODA_E_NCCL, "nccl" #n " not found in libnccl")
    SYM(GetUniqueId);
    SYM(CommInitRank);
    SYM(AllReduce);
    SYM(AllGather);
    SYM(GroupStart);
    SYM(GroupEnd);
    SYM(CommDestroy);
    SYM(GetErrorString);
#undef SYM
    g_nccl.h = h;
    return YODA_OK;
}
#define NC(x)                                                                                      \
    do {                                                                                           \
        ncclResult_t r_ = (x);                                                                     \
        if (r_ != ncclSuccess) return fail(YODA_E_NCCL, "%s failed: %s", #x, g_nccl.GetErrorString(r_)); \
    } while (0)

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
enum { EV_START = 0, EV_PREP, EV_MAX, EV_SCORE, EV_WRITE, EV_H2D0, EV_H2D1, EV_D2H0, EV_D2H1, EV_COUNT };

#ifndef SCORE_NPL
#define SCORE_NPL 2  // nodes per lane in the score pass: 2 measured best on B200 (c3 0.98 ms vs 1.04 at 4, 1.21 at 1; same order at 125k x 10k and 50k x 10k)
#endif

struct yoda_ctx {
    std::mutex mu;
    yoda_config cfg{};
    cudaStream_t own_stream = nullptr, stream = nullptr;
    DevSnap snap{};
    u32 *d_wide_flag = nullptr;  // [2]: snapshot outside the narrow range, node ordering keys changed
    u32 *ord_keys_out = nullptr, *ord_iota = nullptr;  // scratch of the node ordering sort
    void *ord_tmp = nullptr;
    size_t ord_tmp_bytes = 0;
    u32 ord_nodes = 0;           // node count the current DevSnap::order was built for
    bool narrow_ok = false;  // no wide node at all
    bool table_ok = false;   // few enough wide nodes for the table path (they go through the exact kernels, node by node)
    u32 n_wide = 0;          // nodes outside the narrow range (META_WIDE)
    u32 *d_wide_list = nullptr, *d_wide_cnt = nullptr;  // their local indices, gathered at commit
    uint64_t epoch = 0;
    // update staging
    UpdRec *h_upd = nullptr, *d_upd = nullptr;
    u32 upd_cap = 0, n_upd = 0;
    std::vector<int32_t> slot_of;
    bool flushed_uncommitted = false;
    u32 n_nodes_pending = 0;
    // eval workspace
    u32 pods_cap = 0;
    yoda_pod_req *d_req = nullptr, *h_req = nullptr;
    PodC *d_pods = nullptr;
    u64 *d_max6 = nullptr, *d_count = nullptr, *d_hilo = nullptr, *d_ckey = nullptr;
    yoda_pod_result *d_res = nullptr, *h_res = nullptr;
    // result matrices
    size_t raw_cap = 0, norm_cap = 0, bits_cap = 0;  // raw_cap in bytes, the others in elements
    bool res_raw32 = false;  // the raw matrix of the current result holds uint32 (YODA_EVAL_RAW32 honoured)
    i64 *d_raw = nullptr;
    u8 *d_norm = nullptr;
    u32 *d_bits = nullptr;
    u64 stride = 0;
    u32 last_P = 0;
    bool last_full = false;
    // phase state
    u32 ph_P = 0, ph_flags = 0;
    int ph_done = -1;
    // timing
    cudaEvent_t ev[EV_COUNT]{};
    yoda_timing timing{};
    u32 launches = 0;
    // NormalizeScore list scratch
    i64 *d_list = nullptr, *d_lohi = nullptr;
    u32 list_cap = 0;
    // framework tail scratch (yoda_framework_tail)
    u32 *tl_tilecnt = nullptr;
    size_t tl_tilecnt_cap = 0;
    TailOut *tl_out = nullptr;
    u32 tl_out_cap = 0;
    i64 *tl_other = nullptr;
    u32 tl_other_cap = 0;
    // pinned staging of yoda_fetch_rows / yoda_fetch_row (uint32 score rows are widened from here)
    u8 *h_rows = nullptr;
    size_t h_rows_cap = 0;
    // NCCL
    ncclComm_t comm = nullptr;
    // exchanges over peer memory (xchg_kernels.cuh): this shard's arena, the other shards' arenas mapped through CUDA IPC
    bool x_ok = false;             // the peer path is set up on EVERY shard (agreed collectively); otherwise NCCL does the exchanges
    u64 *x_arena = nullptr;
    u32 x_cap = 0;                 // pods per parity the arenas hold
    XPeers x_peers{};
    void *x_opened[X_MAX_RANKS]{};  // IPC mappings to close
    u8 *x_handles = nullptr;       // device scratch of the handle all-gather
    u32 *x_done = nullptr, *x_err = nullptr;
    u64 x_seq = 0;                 // sharded batches so far: the same on every shard
    u32 x_gen[X_PHASES]{};         // exchanges of each phase inside the current batch (a redo repeats phase 0)
    u64 *d_agree = nullptr, *h_agree = nullptr;  // one word all-reduced before a sharded DECIDE score pass (see phase1)
    bool fw_agree = false;                       // this evaluation may ask the other shards which kernels they run
    // fast (narrow) path workspace
    int ph_path = 0;  // path of the evaluation in flight: 0 direct, 1 fast
    u32 fw_pods_cap = 0, fw_htab = 0;
    u64 *fw_gkeys = nullptr, *fw_ckeys = nullptr;
    u32 *fw_gids = nullptr, *fw_cids = nullptr, *fw_crep = nullptr, *fw_slot = nullptr;
    u64 *fw_keys_in = nullptr, *fw_keys_out = nullptr;
    u32 *fw_vals_in = nullptr, *fw_vals_out = nullptr;
    void *fw_cub_tmp = nullptr;
    size_t fw_cub_bytes = 0;
    GroupDesc *fw_gdesc = nullptr;
    PodS *fw_pods = nullptr;
    u32 *fw_lead = nullptr;  // [6][tiles of SCANT pods] leading-run maxima of the maxima scan
    u32 *fw_m_sorted = nullptr, *fw_step = nullptr, *fw_cls_max = nullptr, *fw_cls_order = nullptr, *fw_tables = nullptr;
    u32 fw_tables_classes = 0;
    FastCtl *fw_ctl = nullptr, *fw_hctl = nullptr;
    u32 fw_n_groups = 0, fw_n_classes = 0;
    bool fw_small = false, fw_allow_defer = false, fw_deferred_err = false;
    bool fw_defer_groups = false;  // the group count stays on the device until the class count is read back
    bool fw_gate_ok = false;       // node-sharded FULL batch: fast/exact path selected on the device by FastCtl::err ...
    bool fw_gated = false;         // ... and this evaluation is running that way (both paths launched, see gate_closed)
    bool fw_opt_ok = false;        // the class count may stay on the device too (tables sized by an earlier batch) ...
    bool fw_optimistic = false;    // ... and this evaluation is running that way: no host round trip in the middle of a batch
    bool fw_no_opt = false;        // set while a batch that hit a capacity limit is redone with the counts read back
    bool fw_decide_merged = false;
    bool fw_groups_pending = false;  // the group count (and its capacity check) has not been read back yet
    // sequential-commit scratch
    u64 *sq_alloc = nullptr;
    u32 *sq_stat = nullptr, *sq_pos = nullptr, *sq_order = nullptr, *sq_status = nullptr;
    u32 sq_pods_cap = 0;
    // CUDA graph of the small-batch sequence (<= FAST_SMALL_P pods through yoda_eval_batch: the one-pod scheduling cycle).
    // A cycle is ~10 short kernels, 3 copies and 2 memsets -- launch-bound -- and identical from cycle to cycle until the
    // node count changes, so the second cycle with the same key captures the sequence and later ones replay it.
    struct GraphKey {
        // everything the host decides grids and kernel arguments from; the snapshot's DATA may change (commits), its buffers do not move
        u32 P = 0, flags = 0, n_nodes = 0, n_wide = 0;
        cudaStream_t stream = nullptr;
        bool operator==(const GraphKey &o) const { return P == o.P && flags == o.flags && n_nodes == o.n_nodes && n_wide == o.n_wide && stream == o.stream; }
    };
    struct GraphPost {  // what the host-side state looks like after the sequence was enqueued
        int ph_path = 0, ph_done = -1;
        bool fw_small = false, fw_deferred_err = false, fw_decide_merged = false, res_raw32 = false, last_full = false;
        u32 last_P = 0, fw_n_groups = 0, fw_n_classes = 0, launches = 0;
    };
    GraphKey g_key{}, g_seen{};
    GraphPost g_post{};
    cudaGraphExec_t g_exec = nullptr;
    bool g_capturing = false, g_enabled = true;
    bool g_replayed = false;  // the last evaluation was a graph replay: only its total time was measured
    u32 score_slots = 0, out_slots = 0;  // CTAs of the hot kernels resident on the device at once
    u32 last_score_chunk = 0, last_out_chunk = 0;  // pods per CTA the hot kernels of the last evaluation ran with (yoda_debug_pod_order)
    int n_sms = 0;  // DECIDE on one GPU: the score pass also yields counts and the node choice
};

#define FAST_CLASS_CAP 4096u
#define FAST_TABLE_BYTES (8ull << 30)  // class tables: 32 B per node per class; more classes than fit -> direct path

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = (cudaSetDevice(dev) == cudaSuccess);
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

template <typename T>
static int dev_alloc(T **p, size_t n) {
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T));
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(YODA_E_NOMEM, "cudaMalloc of %zu bytes failed: %s", n * sizeof(T), cudaGetErrorString(e));
    }
    *p = (T *)q;
    return YODA_OK;
}
template <typename T>
static void dev_free(T *&p) {
    if (p) cudaFree(p);
    p = nullptr;
}

static u32 round_up(u32 x, u32 m) { return (x + m - 1) / m * m; }

extern "C" int yoda_abi_version(void) { return YODA_ABI_VERSION; }
extern "C" const char *yoda_last_error(void) { return g_err; }

extern "C" int yoda_create(const yoda_config *cfg, yoda_ctx **out) {
    if (!cfg || !out) return fail(YODA_E_ARG, "yoda_create: null argument");
    if (cfg->max_nodes == 0) return fail(YODA_E_ARG, "yoda_create: max_nodes must be > 0");
    if (cfg->shard_count == 0 || cfg->shard_rank >= cfg->shard_count)
        return fail(YODA_E_ARG, "yoda_create: bad shard rank/count %u/%u", cfg->shard_rank, cfg->shard_count);
    if ((uint64_t)cfg->node_base + cfg->max_nodes > 0xFFFFFFFEull)
        return fail(YODA_E_RANGE, "yoda_create: node_base + max_nodes exceeds 2^32-2");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(YODA_E_CUDA, "no usable CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return fail(YODA_E_ARG, "yoda_create: device %d of %d", cfg->device, ndev);
    DeviceGuard dg(cfg->device);
    if (!dg.ok) return fail(YODA_E_CUDA, "cudaSetDevice(%d) failed", cfg->device);
    yoda_ctx *c = new (std::nothrow) yoda_ctx();
    if (!c) return fail(YODA_E_NOMEM, "out of host memory");
    c->cfg = *cfg;
    auto bail = [&](int rc) {
        yoda_destroy(c);
        return rc;
    };
    if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess)
        return bail(fail(YODA_E_CUDA, "cudaStreamCreate failed"));
    c->stream = c->own_stream;
    {
        const char *g = getenv("YODA_B200_GRAPH");  // YODA_B200_GRAPH=0: never capture
        c->g_enabled = !(g && g[0] == '0');
    }
    if (cudaDeviceGetAttribute(&c->n_sms, cudaDevAttrMultiProcessorCount, cfg->device) != cudaSuccess || c->n_sms <= 0) c->n_sms = 148;
    {
        int per_sm = 0;
        // the score pass keeps its nodes in shared memory: more than the 48 KB a kernel gets without asking
        if (cudaFuncSetAttribute(k_fast_score<SCORE_NPL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScoreSmem<SCORE_NPL, false>)) != cudaSuccess ||
            cudaFuncSetAttribute(k_fast_score<SCORE_NPL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScoreSmem<SCORE_NPL, true>)) != cudaSuccess)
            return bail(fail(YODA_E_CUDA, "cudaFuncSetAttribute(shared memory of the score pass) failed: %s", cudaGetErrorString(cudaGetLastError())));
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fast_score<SCORE_NPL, false>, HOT_THREADS, sizeof(ScoreSmem<SCORE_NPL, false>)) != cudaSuccess || per_sm <= 0) per_sm = 2;
        c->score_slots = (u32)per_sm * (u32)c->n_sms;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fast_out<4, true, false>, HOT_THREADS, 0) != cudaSuccess || per_sm <= 0) per_sm = 2;
        c->out_slots = (u32)per_sm * (u32)c->n_sms;
        cudaGetLastError();
    }
    for (int i = 0; i < EV_COUNT; ++i)
        if (cudaEventCreate(&c->ev[i]) != cudaSuccess) return bail(fail(YODA_E_CUDA, "cudaEventCreate failed"));
    const u32 cap = round_up(cfg->max_nodes, 128);
    DevSnap &s = c->snap;
    s.cap = cap;
    s.n_nodes = 0;
    s.node_base = cfg->node_base;
    int rc = 0;
    for (int f = 0; f < NFIELD && !rc; ++f) rc = dev_alloc(&s.f[f], (size_t)cap * G);
    if (!rc) rc = dev_alloc(&s.card_number, cap);
    if (!rc) rc = dev_alloc(&s.free_sum, cap);
    if (!rc) rc = dev_alloc(&s.total_sum, cap);
    if (!rc) rc = dev_alloc(&s.alloc_sum, cap);
    if (!rc) rc = dev_alloc(&s.stat, cap);
    if (!rc) rc = dev_alloc(&s.meta, cap);
    for (int f = 0; f < NFIELD && !rc; ++f) rc = dev_alloc(&s.n32[f], (size_t)cap * G);
    if (!rc) rc = dev_alloc(&s.cardnum32, cap);
    if (!rc) rc = dev_alloc(&s.stat32, cap);
    if (!rc) rc = dev_alloc(&c->d_wide_flag, 2);
    if (!rc) rc = dev_alloc(&s.tk32, (size_t)cap * G);
    if (!rc) rc = dev_alloc(&s.order, cap);
    if (!rc) rc = dev_alloc(&s.order_key, cap);
    if (!rc) rc = dev_alloc(&c->ord_keys_out, cap);
    if (!rc) rc = dev_alloc(&c->ord_iota, cap);
    if (!rc) {
        if (cub::DeviceRadixSort::SortPairs(nullptr, c->ord_tmp_bytes, s.order_key, c->ord_keys_out, c->ord_iota, s.order, (int)cap, 0, 32,
                                            c->stream) != cudaSuccess ||
            cudaMalloc(&c->ord_tmp, std::max<size_t>(c->ord_tmp_bytes, 16)) != cudaSuccess)
            rc = fail(YODA_E_NOMEM, "scratch of the node ordering sort");
    }
    if (!rc) rc = dev_alloc(&c->d_lohi, 2);
    if (rc) return bail(rc);
    for (int f = 0; f < NFIELD; ++f) cudaMemsetAsync(s.f[f], 0, (size_t)cap * G * 8, c->stream);
    cudaMemsetAsync(s.card_number, 0, (size_t)cap * 8, c->stream);
    cudaMemsetAsync(s.free_sum, 0, (size_t)cap * 8, c->stream);
    cudaMemsetAsync(s.total_sum, 0, (size_t)cap * 8, c->stream);
    cudaMemsetAsync(s.alloc_sum, 0, (size_t)cap * 8, c->stream);
    cudaMemsetAsync(s.stat, 0, (size_t)cap * 8, c->stream);
    cudaMemsetAsync(s.meta, 0, (size_t)cap * 4, c->stream);
    cudaMemsetAsync(s.order_key, 0xFF, (size_t)cap * 4, c->stream);  // no node has this key: the first commit orders
    // derived planes of nodes no update ever touched: what the derive step would compute for "no SCV"
    cudaMemsetAsync(s.stat32, 0, (size_t)cap * 4, c->stream);
    cudaMemsetAsync(s.cardnum32, 0, (size_t)cap * 4, c->stream);
    cudaMemsetAsync(s.tk32, 0, (size_t)cap * G * 4, c->stream);
    for (int f = 0; f < NFIELD; ++f) cudaMemsetAsync(s.n32[f], 0, (size_t)cap * G * 4, c->stream);
    cudaMemsetAsync(c->d_wide_flag, 0, 2 * sizeof(u32), c->stream);
    c->upd_cap = std::min<u32>(cfg->max_nodes, 1u << 16);
    if (cudaMallocHost((void **)&c->h_upd, (size_t)c->upd_cap * sizeof(UpdRec)) != cudaSuccess)
        return bail(fail(YODA_E_NOMEM, "cudaMallocHost of the update staging buffer failed"));
    rc = dev_alloc(&c->d_upd, c->upd_cap);
    if (rc) return bail(rc);
    c->slot_of.assign(cfg->max_nodes, -1);
    if (cudaStreamSynchronize(c->stream) != cudaSuccess) return bail(fail(YODA_E_CUDA, "initial memset failed"));
    *out = c;
    return YODA_OK;
}

static int nccl_host_barrier(yoda_ctx *c, u64 mine, int op_min, u64 *result);
static void xchg_release(yoda_ctx *c);

extern "C" void yoda_destroy(yoda_ctx *c) {
    if (!c) return;
    DeviceGuard dg(c->cfg.device);
    cudaDeviceSynchronize();
    if (c->comm && c->x_ok) {
        // a peer may still be reading this shard's arena: leave together (yoda_destroy of a sharded context is collective)
        u64 dummy;
        nccl_host_barrier(c, 0, 0, &dummy);
    }
    xchg_release(c);
    dev_free(c->x_handles); dev_free(c->x_done); dev_free(c->x_err);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    dev_free(c->d_agree);
    if (c->h_agree) cudaFreeHost(c->h_agree);
    DevSnap &s = c->snap;
    for (int f = 0; f < NFIELD; ++f) dev_free(s.f[f]);
    dev_free(s.card_number); dev_free(s.free_sum); dev_free(s.total_sum); dev_free(s.alloc_sum);
    dev_free(s.stat); dev_free(s.meta);
    for (int f = 0; f < NFIELD; ++f) dev_free(s.n32[f]);
    dev_free(s.cardnum32); dev_free(s.stat32); dev_free(s.tk32); dev_free(s.order); dev_free(s.order_key);
    dev_free(c->ord_keys_out); dev_free(c->ord_iota);
    if (c->ord_tmp) cudaFree(c->ord_tmp);
    if (c->g_exec) cudaGraphExecDestroy(c->g_exec);
    dev_free(c->d_wide_list); dev_free(c->d_wide_cnt);
    dev_free(c->tl_tilecnt); dev_free(c->tl_out); dev_free(c->tl_other);
    dev_free(c->d_wide_flag); dev_free(c->d_lohi); dev_free(c->d_list); dev_free(c->d_upd);
    dev_free(c->d_req); dev_free(c->d_pods); dev_free(c->d_max6); dev_free(c->d_count);
    dev_free(c->d_hilo); dev_free(c->d_ckey); dev_free(c->d_res);
    dev_free(c->d_raw); dev_free(c->d_norm); dev_free(c->d_bits);
    dev_free(c->fw_gkeys); dev_free(c->fw_ckeys); dev_free(c->fw_gids); dev_free(c->fw_cids); dev_free(c->fw_crep);
    dev_free(c->fw_slot); dev_free(c->fw_keys_in); dev_free(c->fw_keys_out);
    dev_free(c->fw_vals_in); dev_free(c->fw_vals_out); dev_free(c->fw_gdesc); dev_free(c->fw_pods);
    dev_free(c->fw_lead);
    dev_free(c->fw_m_sorted); dev_free(c->fw_step); dev_free(c->fw_cls_max); dev_free(c->fw_cls_order); dev_free(c->fw_tables); dev_free(c->fw_ctl);
    dev_free(c->sq_alloc); dev_free(c->sq_stat); dev_free(c->sq_pos); dev_free(c->sq_order); dev_free(c->sq_status);
    if (c->fw_cub_tmp) cudaFree(c->fw_cub_tmp);
    if (c->fw_hctl) cudaFreeHost(c->fw_hctl);
    if (c->h_upd) cudaFreeHost(c->h_upd);
    if (c->h_rows) cudaFreeHost(c->h_rows);
    if (c->h_req) cudaFreeHost(c->h_req);
    if (c->h_res) cudaFreeHost(c->h_res);
    for (int i = 0; i < EV_COUNT; ++i)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

extern "C" int yoda_set_stream(yoda_ctx *c, void *cuda_stream) {
    if (!c) return fail(YODA_E_ARG, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    CU(cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return YODA_OK;
}

// ------------------------------------------------------------------------------------------
// snapshot maintenance
// ------------------------------------------------------------------------------------------
static int flush_updates(yoda_ctx *c) {
    if (c->n_upd == 0) return YODA_OK;
    CU(cudaMemcpyAsync(c->d_upd, c->h_upd, (size_t)c->n_upd * sizeof(UpdRec), cudaMemcpyHostToDevice, c->stream));
    const u32 threads = c->n_upd * G;
    k_scatter_updates<<<(threads + 255) / 256, 256, 0, c->stream>>>(c->snap, c->d_upd, c->n_upd);
    CU(cudaGetLastError());
    k_prepare_updates<<<(c->n_upd + 127) / 128, 128, 0, c->stream>>>(c->snap, c->d_upd, c->n_upd, c->d_wide_flag);  // derived data of the touched nodes
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(c->stream));  // the pinned buffer is reused right away
    for (u32 i = 0; i < c->n_upd; ++i) c->slot_of[c->h_upd[i].idx] = -1;
    c->n_upd = 0;
    return YODA_OK;
}

static int stage_slot(yoda_ctx *c, u32 idx, UpdRec **out) {
    if (idx >= c->cfg.max_nodes) return fail(YODA_E_RANGE, "node index %u >= max_nodes %u", idx, c->cfg.max_nodes);
    int32_t sl = c->slot_of[idx];
    if (sl < 0) {
        if (c->n_upd == c->upd_cap) {
            TRY(flush_updates(c));
            c->flushed_uncommitted = true;
        }
        sl = (int32_t)c->n_upd++;
        c->slot_of[idx] = sl;
        c->h_upd[sl].idx = idx;
        c->h_upd[sl].op = 0;
        c->h_upd[sl].alloc = 0;
    }
    c->n_nodes_pending = std::max(c->n_nodes_pending, idx + 1);
    *out = &c->h_upd[sl];
    return YODA_OK;
}

static int check_record(const yoda_node_record *r, u32 idx) {
    if (r->n_cards > YODA_MAX_CARDS)
        return fail(YODA_E_RANGE, "node %u: %u cards, the device snapshot holds at most %d per node", idx, r->n_cards,
                    YODA_MAX_CARDS);
    if (r->total_memory_sum == 0)
        return fail(YODA_E_DIV0, "node %u: TotalMemorySum == 0 (the reference divides by it, algorithm.go:71,86)", idx);
    return YODA_OK;
}

static int upsert_locked(yoda_ctx *c, u32 idx, const yoda_node_record *rec) {
    TRY(check_record(rec, idx));
    UpdRec *u;
    TRY(stage_slot(c, idx, &u));
    u->rec = *rec;
    u->op = (u->op | UPD_REC) & ~UPD_REMOVE;
    return YODA_OK;
}

extern "C" int yoda_node_upsert(yoda_ctx *c, u32 idx, const yoda_node_record *rec) {
    if (!c || !rec) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    return upsert_locked(c, idx, rec);
}

extern "C" int yoda_node_upsert_many(yoda_ctx *c, u32 n, const u32 *idx, const yoda_node_record *recs) {
    if (!c || (n && (!idx || !recs))) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    for (u32 i = 0; i < n; ++i) TRY(upsert_locked(c, idx[i], &recs[i]));
    return YODA_OK;
}

extern "C" int yoda_node_set_alloc(yoda_ctx *c, u32 idx, uint64_t a) {
    if (!c) return fail(YODA_E_ARG, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    UpdRec *u;
    TRY(stage_slot(c, idx, &u));
    u->alloc = a;
    u->op |= UPD_ALLOC;
    return YODA_OK;
}

extern "C" int yoda_node_set_alloc_many(yoda_ctx *c, u32 n, const u32 *idx, const uint64_t *a) {
    if (!c || (n && (!idx || !a))) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    for (u32 i = 0; i < n; ++i) {
        UpdRec *u;
        TRY(stage_slot(c, idx[i], &u));
        u->alloc = a[i];
        u->op |= UPD_ALLOC;
    }
    return YODA_OK;
}

extern "C" int yoda_node_remove(yoda_ctx *c, u32 idx) {
    if (!c) return fail(YODA_E_ARG, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    UpdRec *u;
    TRY(stage_slot(c, idx, &u));
    u->op = (u->op | UPD_REMOVE) & ~UPD_REC;
    return YODA_OK;
}

extern "C" int yoda_commit(yoda_ctx *c, uint64_t *epoch) {
    if (!c) return fail(YODA_E_ARG, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    TRY(flush_updates(c));
    c->snap.n_nodes = c->n_nodes_pending;
    c->flushed_uncommitted = false;
    const u32 N = c->snap.n_nodes;
    // the derived per-node data was brought up to date record by record (flush_updates); what is left is the
    // snapshot-wide verdict: how many nodes are outside the narrow range, did any ordering key change
    u32 flag[2] = {0, 0};
    CU(cudaMemcpyAsync(flag, c->d_wide_flag, sizeof flag, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemsetAsync(c->d_wide_flag + 1, 0, sizeof(u32), c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->n_wide = flag[0];
    c->narrow_ok = (c->n_wide == 0);
    // A few wide nodes (one misreported SCV field, a CardNumber beyond 32 bits) do not cost the batch the table path:
    // the tables skip them and the exact kernels evaluate just those nodes.  Mostly-wide snapshots go direct wholesale.
    c->table_ok = c->n_wide <= std::max(64u, N / 8);
    if (c->n_wide && c->table_ok) {
        if (!c->d_wide_list) {
            TRY(dev_alloc(&c->d_wide_list, c->snap.cap));
            TRY(dev_alloc(&c->d_wide_cnt, 1));
        }
        CU(cudaMemsetAsync(c->d_wide_cnt, 0, sizeof(u32), c->stream));
        k_wide_list<<<(N + 255) / 256, 256, 0, c->stream>>>(c->snap.meta, N, c->d_wide_list, c->d_wide_cnt, c->snap.cap);
        CU(cudaGetLastError());
    }
    if (N && (flag[1] || N != c->ord_nodes)) {  // metric ticks that leave the clocks alone keep the order
        k_iota<<<(N + 255) / 256, 256, 0, c->stream>>>(c->ord_iota, N);
        CU(cudaGetLastError());
        size_t bytes = c->ord_tmp_bytes;
        CU(cub::DeviceRadixSort::SortPairs(c->ord_tmp, bytes, c->snap.order_key, c->ord_keys_out, c->ord_iota, c->snap.order, (int)N, 0, 32,
                                           c->stream));
        c->ord_nodes = N;
    }
    c->epoch++;
    c->last_P = 0;  // result matrices belong to the previous snapshot
    if (epoch) *epoch = c->epoch;
    return YODA_OK;
}

extern "C" int yoda_node_count(yoda_ctx *c, u32 *n) {
    if (!c || !n) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    *n = c->snap.n_nodes;
    return YODA_OK;
}

// ------------------------------------------------------------------------------------------
// evaluation
// ------------------------------------------------------------------------------------------
// The captured small-batch graph holds device addresses: whenever a buffer it may reference is reallocated, it goes.
static void graph_drop(yoda_ctx *c) {
    if (c->g_exec) cudaGraphExecDestroy(c->g_exec);
    c->g_exec = nullptr;
    c->g_key = yoda_ctx::GraphKey{};
    c->g_seen = yoda_ctx::GraphKey{};
}

static int ensure_pods(yoda_ctx *c, u32 P) {
    if (P <= c->pods_cap) return YODA_OK;
    const u32 cap = round_up(std::max(P, 1024u), 1024);
    CU(cudaStreamSynchronize(c->stream));
    graph_drop(c);
    dev_free(c->d_req); dev_free(c->d_pods); dev_free(c->d_max6); dev_free(c->d_count);
    dev_free(c->d_hilo); dev_free(c->d_ckey); dev_free(c->d_res);
    if (c->h_req) cudaFreeHost(c->h_req);
    if (c->h_res) cudaFreeHost(c->h_res);
    c->h_req = nullptr;
    c->h_res = nullptr;
    c->pods_cap = 0;
    TRY(dev_alloc(&c->d_req, cap));
    TRY(dev_alloc(&c->d_pods, cap));
    TRY(dev_alloc(&c->d_max6, (size_t)cap * 6));
    TRY(dev_alloc(&c->d_count, cap));
    TRY(dev_alloc(&c->d_hilo, (size_t)cap * 2));
    TRY(dev_alloc(&c->d_ckey, cap));
    TRY(dev_alloc(&c->d_res, cap));
    if (cudaMallocHost((void **)&c->h_req, (size_t)cap * sizeof(yoda_pod_req)) != cudaSuccess ||
        cudaMallocHost((void **)&c->h_res, (size_t)cap * sizeof(yoda_pod_result)) != cudaSuccess)
        return fail(YODA_E_NOMEM, "cudaMallocHost of the pod staging buffers failed");
    c->pods_cap = cap;
    return YODA_OK;
}

static int ensure_raw(yoda_ctx *c, size_t bytes) {
    if (bytes <= c->raw_cap) return YODA_OK;
    CU(cudaStreamSynchronize(c->stream));
    graph_drop(c);
    dev_free(c->d_raw);
    c->raw_cap = 0;
    TRY(dev_alloc((u8 **)&c->d_raw, bytes));
    c->raw_cap = bytes;
    return YODA_OK;
}

static int ensure_matrices(yoda_ctx *c, u32 P, u64 stride, u32 raw_elem_bytes) {
    const size_t need = (size_t)P * stride, need_bits = (size_t)P * (stride / 32);
    TRY(ensure_raw(c, need * raw_elem_bytes));
    if (need > c->norm_cap) {
        CU(cudaStreamSynchronize(c->stream));
        graph_drop(c);
        dev_free(c->d_norm);
        c->norm_cap = 0;
        TRY(dev_alloc(&c->d_norm, need));
        c->norm_cap = need;
    }
    if (need_bits > c->bits_cap) {
        CU(cudaStreamSynchronize(c->stream));
        graph_drop(c);
        dev_free(c->d_bits);
        c->bits_cap = 0;
        TRY(dev_alloc(&c->d_bits, need_bits));
        c->bits_cap = need_bits;
    }
    return YODA_OK;
}

// phase events; not while the small-batch sequence is being captured into a graph (see yoda_eval_batch)
static int ev_record(yoda_ctx *c, int which) {
    if (c->g_capturing) return YODA_OK;
    CU(cudaEventRecord(c->ev[which], c->stream));
    return YODA_OK;
}

#define LAUNCHED(c)          \
    do {                     \
        CU(cudaGetLastError()); \
        (c)->launches++;     \
    } while (0)

// ---- fast path workspace ------------------------------------------------------------------
static u32 fast_class_cap(const yoda_ctx *c) {
    const u64 per_class = (u64)c->snap.cap * G * sizeof(u32);
    return (u32)std::max<u64>(1, std::min<u64>(FAST_CLASS_CAP, FAST_TABLE_BYTES / per_class));
}

static int ensure_fast(yoda_ctx *c, u32 P) {
    if (!c->fw_ctl) {
        TRY(dev_alloc(&c->fw_ctl, 1));
        if (cudaMallocHost((void **)&c->fw_hctl, sizeof(FastCtl)) != cudaSuccess) return fail(YODA_E_NOMEM, "cudaMallocHost failed");
        TRY(dev_alloc(&c->fw_gdesc, FAST_GROUP_CAP));
        TRY(dev_alloc(&c->fw_cls_max, (size_t)FAST_CLASS_CAP * 6));
        TRY(dev_alloc(&c->fw_cls_order, (size_t)FAST_CLASS_CAP));
    }
    if (P <= c->fw_pods_cap) return YODA_OK;
    CU(cudaStreamSynchronize(c->stream));
    graph_drop(c);
    const u32 cap = c->pods_cap;  // grown by ensure_pods just before
    u32 htab = 1024;
    while (htab < 2 * cap) htab <<= 1;
    dev_free(c->fw_gkeys); dev_free(c->fw_ckeys); dev_free(c->fw_gids); dev_free(c->fw_cids); dev_free(c->fw_crep);
    dev_free(c->fw_slot); dev_free(c->fw_keys_in); dev_free(c->fw_keys_out);
    dev_free(c->fw_vals_in); dev_free(c->fw_vals_out); dev_free(c->fw_pods);
    dev_free(c->fw_m_sorted); dev_free(c->fw_step); dev_free(c->fw_lead);
    if (c->fw_cub_tmp) cudaFree(c->fw_cub_tmp);
    c->fw_cub_tmp = nullptr;
    c->fw_pods_cap = 0;
    TRY(dev_alloc(&c->fw_gkeys, htab));
    TRY(dev_alloc(&c->fw_ckeys, htab));
    TRY(dev_alloc(&c->fw_gids, htab));
    TRY(dev_alloc(&c->fw_cids, htab));
    TRY(dev_alloc(&c->fw_crep, htab));
    TRY(dev_alloc(&c->fw_slot, cap));
    TRY(dev_alloc(&c->fw_keys_in, cap));
    TRY(dev_alloc(&c->fw_keys_out, cap));
    TRY(dev_alloc(&c->fw_vals_in, cap));
    TRY(dev_alloc(&c->fw_vals_out, cap));
    TRY(dev_alloc(&c->fw_pods, cap));
    TRY(dev_alloc(&c->fw_m_sorted, cap));
    TRY(dev_alloc(&c->fw_step, (size_t)cap * 6));
    TRY(dev_alloc(&c->fw_lead, ((size_t)cap / SCANT + 1) * 6));
    size_t bytes = 0;
    CU(cub::DeviceRadixSort::SortPairs(nullptr, bytes, c->fw_keys_in, c->fw_keys_out, c->fw_vals_in, c->fw_vals_out, (int)cap, 0,
                                       33 + FAST_GID_BITS, c->stream));
    if (cudaMalloc(&c->fw_cub_tmp, std::max<size_t>(bytes, 16)) != cudaSuccess) return fail(YODA_E_NOMEM, "sort scratch allocation failed");
    c->fw_cub_bytes = bytes;
    c->fw_htab = htab;
    c->fw_pods_cap = cap;
    return YODA_OK;
}

static int fast_read_ctl(yoda_ctx *c) {
    CU(cudaMemcpyAsync(c->fw_hctl, c->fw_ctl, sizeof(FastCtl), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return YODA_OK;
}

// ---- direct path bodies ---------------------------------------------------------------------
static int direct_pods_launch(yoda_ctx *c, int mode, bool gated = false) {
    const u32 P = c->ph_P, N = c->snap.n_nodes;
    if (!N) return YODA_OK;
    const u32 gx = (P + DPT - 1) / DPT;
    u32 chunks = std::max(1u, std::min((N + DTN - 1) / DTN, (592u + gx - 1) / gx));
    chunks = std::min(chunks, 65535u);
    const u32 npc = round_up((N + chunks - 1) / chunks, DTN);
    dim3 grid(gx, (N + npc - 1) / npc);
    DirectOut none{};
    if (gated) {
        none.gate = &c->fw_ctl->err;
        none.gate_mode = GATE_IF_SET;
    }
    if (mode == 0)
        k_direct_pods<0><<<grid, DPT, 0, c->stream>>>(c->snap, c->d_pods, P, npc, c->d_max6, c->d_hilo, nullptr, 0, none);
    else
        k_direct_pods<1><<<grid, DPT, 0, c->stream>>>(c->snap, c->d_pods, P, npc, c->d_max6, c->d_hilo, nullptr, 0, none);
    LAUNCHED(c);
    return YODA_OK;
}

// The wide nodes of a mixed snapshot, evaluated exactly while the table path takes all other nodes.
// mode 0: their maxima, 1: their lowest/highest, 2: their part of the output pass.  thread = pod.
static int wide_pods_launch(yoda_ctx *c, int mode) {
    const u32 P = c->ph_P, W = c->n_wide;
    if (!W || !c->snap.n_nodes) return YODA_OK;
    const u32 gx = (P + DPT - 1) / DPT;
    u32 chunks = std::max(1u, std::min((W + DTN - 1) / DTN, (592u + gx - 1) / gx));
    chunks = std::min(chunks, 65535u);
    const u32 npc = round_up((W + chunks - 1) / chunks, DTN);
    dim3 grid(gx, (W + npc - 1) / npc);
    DirectOut out{};
    if (c->fw_gated) {
        out.gate = &c->fw_ctl->err;
        out.gate_mode = GATE_IF_ZERO;
    }
    if (mode == 0) {
        k_direct_pods<0><<<grid, DPT, 0, c->stream>>>(c->snap, c->d_pods, P, npc, c->d_max6, c->d_hilo, c->d_wide_list, W, out);
    } else if (mode == 1) {
        k_direct_pods<1><<<grid, DPT, 0, c->stream>>>(c->snap, c->d_pods, P, npc, c->d_max6, c->d_hilo, c->d_wide_list, W, out);
    } else {
        const bool full = (c->ph_flags & YODA_EVAL_FULL) != 0;
        out.count = c->d_count;
        out.ckey = c->d_ckey;
        out.raw = full ? c->d_raw : nullptr;
        out.norm = full ? c->d_norm : nullptr;
        out.bits = full ? c->d_bits : nullptr;
        out.stride = c->stride;
        out.packed = c->fw_decide_merged ? 1u : 0u;
        if (c->fw_decide_merged) {
            k_wide_promote<<<(P + 255) / 256, 256, 0, c->stream>>>(c->d_hilo, c->d_count, c->d_ckey, P);
            LAUNCHED(c);
        }
        k_direct_pods<2><<<grid, DPT, 0, c->stream>>>(c->snap, c->d_pods, P, npc, c->d_max6, c->d_hilo, c->d_wide_list, W, out);
    }
    LAUNCHED(c);
    return YODA_OK;
}

static int direct_out_launch(yoda_ctx *c, bool gated = false) {
    const u32 P = c->ph_P, N = c->snap.n_nodes;
    if (!N) return YODA_OK;
    cudaStream_t st = c->stream;
    const bool full = (c->ph_flags & YODA_EVAL_FULL) != 0;
    const u32 gx = (N + DCT - 1) / DCT;
    const u32 want_y = std::max(1u, 2368u / gx);
    u32 ppc = std::max(1u, (P + want_y - 1) / want_y);
    if (!gated) ppc = std::min(ppc, 64u);  // (a gated launch usually finds its gate closed: one wave of CTAs, not thousands)
    ppc = std::max(ppc, (P + 65534u) / 65535u);
    dim3 grid(gx, (P + ppc - 1) / ppc);
    if (full && !gated) {
        TRY(ensure_raw(c, (size_t)P * c->stride * sizeof(i64)));  // a YODA_EVAL_RAW32 batch that fell back to these kernels
        c->res_raw32 = false;
    }
    const u32 *gate = gated ? &c->fw_ctl->err : nullptr;
    const int gm = gated ? GATE_IF_SET : GATE_NONE;
    if (full)
        k_direct_out<true><<<grid, DCT, 0, st>>>(c->snap, c->d_pods, P, ppc, c->d_max6, c->d_count, c->d_hilo, c->d_ckey, c->d_raw,
                                                 c->d_norm, c->d_bits, c->stride, gate, gm);
    else
        k_direct_out<false><<<grid, DCT, 0, st>>>(c->snap, c->d_pods, P, ppc, c->d_max6, c->d_count, c->d_hilo, c->d_ckey, nullptr,
                                                  nullptr, nullptr, c->stride, gate, gm);
    LAUNCHED(c);
    return YODA_OK;
}

// ---- fast path bodies -------------------------------------------------------------------------
// groups -> sorted pods -> maxima.  Returns with ph_path = 0 if the batch has too many groups.
// step planes -> per-pod maxima (two launches, see k_max_scan)
static int max_scan_launch(yoda_ctx *c, u32 P) {
    const u32 n_tiles = (P + SCANT - 1) / SCANT;
    k_max_tile_lead<<<dim3(n_tiles, 6), SCANT, 0, c->stream>>>(c->fw_gdesc, c->fw_step, c->fw_pods, c->fw_lead, P, n_tiles);
    LAUNCHED(c);
    k_max_scan<<<dim3(n_tiles, 6), SCANT, 0, c->stream>>>(c->fw_gdesc, c->fw_step, c->fw_pods, c->fw_lead, c->d_max6, P, n_tiles);
    LAUNCHED(c);
    return YODA_OK;
}

static int fast_phase0(yoda_ctx *c) {
    const u32 P = c->ph_P, N = c->snap.n_nodes;
    cudaStream_t st = c->stream;
    const u32 pb = (P + 255) / 256;
    if (!c->fw_small) CU(cudaMemsetAsync(c->fw_ctl, 0, sizeof(FastCtl), st));
    if (c->fw_small) {  // one pod = one group = one class: no hashing, no sort, no host round trip; set up by k_small_setup (phase0)
        c->fw_n_groups = P;
        TRY(ev_record(c, EV_PREP));
        if (N) {
            k_max_thresholds<<<dim3((N + MAXT - 1) / MAXT, 1), MAXT, 0, st>>>(c->snap, c->fw_gdesc, P, nullptr, c->fw_m_sorted, c->fw_step, P);
            LAUNCHED(c);
        }
        k_small_max_out<<<1, FAST_SMALL_P * 6, 0, st>>>(c->fw_step, c->fw_pods, c->d_max6, P);
        LAUNCHED(c);
        return YODA_OK;
    }
    k_hash_clear<<<(c->fw_htab + 255) / 256, 256, 0, st>>>(c->fw_gkeys, c->fw_gids, c->fw_htab);
    LAUNCHED(c);
    k_group_insert<<<(pb + INSERT_PER - 1) / INSERT_PER, 256, 0, st>>>(c->d_pods, P, c->fw_gkeys, c->fw_htab - 1, c->fw_slot);
    LAUNCHED(c);
    k_hash_compact<<<(c->fw_htab + 255) / 256, 256, 0, st>>>(c->fw_gkeys, c->fw_gids, c->fw_htab, &c->fw_ctl->n_groups);
    LAUNCHED(c);
    // On one GPU the group count stays on the device until the class count is read back (fast_classes):
    // the kernels below take it from ctl.  Sharded / phase-split evaluations read it now, because the
    // maxima are exchanged right after this phase and a fallback decided later could not undo that.
    const bool count_on_device = c->fw_defer_groups;
    u32 groups_bound = std::min(P, FAST_GROUP_CAP);
    if (!count_on_device) {
        TRY(fast_read_ctl(c));
        c->fw_n_groups = c->fw_hctl->n_groups;
        if (c->fw_n_groups > FAST_GROUP_CAP) {
            c->ph_path = 0;
            return YODA_OK;
        }
        groups_bound = c->fw_n_groups;
    }
    const FastCtl *dctl = count_on_device ? c->fw_ctl : nullptr;
    k_group_assign<<<pb, 256, 0, st>>>(c->d_pods, P, c->fw_slot, c->fw_gids, c->fw_gdesc, FAST_GROUP_CAP, c->fw_keys_in, c->fw_vals_in,
                                       c->fw_ctl);
    LAUNCHED(c);
    size_t bytes = c->fw_cub_bytes;
    CU(cub::DeviceRadixSort::SortPairs(c->fw_cub_tmp, bytes, c->fw_keys_in, c->fw_keys_out, c->fw_vals_in, c->fw_vals_out, (int)P, 0,
                                       33 + FAST_GID_BITS, st));
    c->launches += 6;  // cub radix sort passes (library kernels)
    k_build_sorted<<<pb, 256, 0, st>>>(c->d_pods, P, c->fw_keys_out, c->fw_vals_out, c->fw_gdesc, FAST_GROUP_CAP, c->fw_pods,
                                       c->fw_m_sorted);
    LAUNCHED(c);
    TRY(ev_record(c, EV_PREP));
    CU(cudaMemsetAsync(c->fw_step, 0, (size_t)P * 6 * sizeof(u32), st));
    if (N) {
        // eight groups per CTA pass: the per-node setup (loads, k-th largest FreeMemory) costs about as much as one group
        const dim3 grid((N + MAXT - 1) / MAXT, std::max(1u, std::min(32u, (groups_bound + MAXG_TILE - 1) / MAXG_TILE)));
        k_max_thresholds<<<grid, MAXT, 0, st>>>(c->snap, c->fw_gdesc, groups_bound, dctl, c->fw_m_sorted, c->fw_step, P);
        LAUNCHED(c);
    }
    TRY(max_scan_launch(c, P));
    c->fw_groups_pending = count_on_device;
    return YODA_OK;
}

// Pods per CTA of the two hot kernels.  The maximum for big batches; for small and medium ones the last wave of CTAs
// would be mostly empty, so the chunk shrinks until the CTAs fill whole waves (slots = CTAs resident at once).
static u32 hot_chunk(u32 n_pods, u32 tiles, u32 slots, u32 max_chunk) {
    const u32 chunks = (n_pods + max_chunk - 1) / max_chunk;
    const u32 waves = (tiles * chunks + slots - 1) / slots;
    const u32 want = std::max(chunks, waves * slots / std::max(tiles, 1u));  // chunks that still fit those waves
    u32 chunk = (n_pods + want - 1) / want;
    chunk = std::min<u32>(max_chunk, std::max<u32>(16u, (chunk + SCORE_BLK - 1) / SCORE_BLK * SCORE_BLK));
    return chunk;
}

static int fast_score_launch(yoda_ctx *c, u32 p_begin, u32 p_end) {
    if (p_end <= p_begin || !c->snap.n_nodes) return YODA_OK;
    const u32 tiles = (c->snap.n_nodes + HOT_WARPS * SCORE_NPL * 32 - 1) / (HOT_WARPS * SCORE_NPL * 32);
    const u32 chunk = hot_chunk(p_end - p_begin, tiles, c->score_slots, SCORE_CHUNK);
    c->last_score_chunk = chunk;
    const dim3 grid(tiles, (p_end - p_begin + chunk - 1) / chunk);
    const u32 *gate = c->fw_gated ? &c->fw_ctl->err : nullptr;
    if (c->fw_decide_merged)
        k_fast_score<SCORE_NPL, true><<<grid, HOT_THREADS, sizeof(ScoreSmem<SCORE_NPL, true>), c->stream>>>(c->snap, c->fw_pods, p_begin, p_end, c->fw_tables, c->d_hilo,
                                                                           c->d_count, c->d_ckey, chunk, gate);
    else
        k_fast_score<SCORE_NPL, false><<<grid, HOT_THREADS, sizeof(ScoreSmem<SCORE_NPL, false>), c->stream>>>(c->snap, c->fw_pods, p_begin, p_end, c->fw_tables, c->d_hilo,
                                                                            c->d_count, c->d_ckey, chunk, gate);
    LAUNCHED(c);
    return YODA_OK;
}

static int fast_out_launch(yoda_ctx *c, u32 p_begin, u32 p_end) {
    if (p_end <= p_begin || !c->snap.n_nodes) return YODA_OK;
    constexpr int NPL = 4;
    const u32 tiles = (c->snap.n_nodes + HOT_WARPS * NPL * 32 - 1) / (HOT_WARPS * NPL * 32);
    const u32 chunk = hot_chunk(p_end - p_begin, tiles, c->out_slots, HOT_CHUNK);
    c->last_out_chunk = chunk;
    const dim3 grid(tiles, (p_end - p_begin + chunk - 1) / chunk);
    const bool full = (c->ph_flags & YODA_EVAL_FULL) != 0;
    const u32 *gate = c->fw_gated ? &c->fw_ctl->err : nullptr;
    if (full) c->res_raw32 = (c->ph_flags & YODA_EVAL_RAW32) != 0 && c->narrow_ok;  // a wide node's Score may not fit 32 bits
    if (full && c->res_raw32)
        k_fast_out<4, true, true><<<grid, HOT_THREADS, 0, c->stream>>>(c->snap, c->fw_pods, c->d_hilo, p_begin, p_end, c->fw_tables, c->d_count,
                                                                       c->d_ckey, c->d_raw, c->d_norm, c->d_bits, c->stride, chunk, gate);
    else if (full)
        k_fast_out<4, true, false><<<grid, HOT_THREADS, 0, c->stream>>>(c->snap, c->fw_pods, c->d_hilo, p_begin, p_end, c->fw_tables, c->d_count,
                                                                        c->d_ckey, c->d_raw, c->d_norm, c->d_bits, c->stride, chunk, gate);
    else
        k_fast_out<4, false, false><<<grid, HOT_THREADS, 0, c->stream>>>(c->snap, c->fw_pods, c->d_hilo, p_begin, p_end, c->fw_tables, c->d_count,
                                                                  c->d_ckey, nullptr, nullptr, nullptr, c->stride, chunk, gate);
    LAUNCHED(c);
    return YODA_OK;
}

static int exchange_locked(yoda_ctx *c, int phase);

// classes -> tables.  Returns with ph_path = 0 if the batch cannot use the tables.
static int fast_classes(yoda_ctx *c) {
    const u32 P = c->ph_P, N = c->snap.n_nodes;
    cudaStream_t st = c->stream;
    const u32 pb = (P + 255) / 256;
    const u32 ccap = fast_class_cap(c);
    if (c->fw_small) {
        k_small_classes<<<1, 64, 0, st>>>(P, c->d_max6, c->fw_cls_max, c->fw_ctl);
        LAUNCHED(c);
        c->fw_n_classes = P;
        if (c->fw_n_classes > c->fw_tables_classes) {
            CU(cudaStreamSynchronize(st));
            graph_drop(c);
            dev_free(c->fw_tables);
            c->fw_tables_classes = 0;
            const u32 want = std::min(ccap, 64u);
            TRY(dev_alloc(&c->fw_tables, (size_t)want * c->snap.cap * G));
            c->fw_tables_classes = want;
        }
        if (N) {
            k_class_tables<<<dim3((N + CLST - 1) / CLST, (P + CLS_PER - 1) / CLS_PER), CLST, 0, st>>>(c->snap, c->fw_cls_max, P, c->fw_tables, c->fw_ctl, false, nullptr);
            LAUNCHED(c);
        }
        if (c->fw_allow_defer) {
            c->fw_deferred_err = true;  // checked once at the end of the evaluation instead of stalling here
            return YODA_OK;
        }
        if (c->fw_gate_ok) {
            c->fw_gated = true;  // both paths are launched, the device picks (gate_closed)
            return YODA_OK;
        }
        TRY(fast_read_ctl(c));
        if (c->fw_hctl->err) c->ph_path = 0;
        return YODA_OK;
    }
    k_hash_clear<<<(c->fw_htab + 255) / 256, 256, 0, st>>>(c->fw_ckeys, c->fw_crep, c->fw_htab);
    LAUNCHED(c);
    k_class_insert<<<pb, 256, 0, st>>>(c->fw_pods, P, c->d_max6, c->fw_ckeys, c->fw_htab - 1, c->fw_crep, c->fw_slot);
    LAUNCHED(c);
    k_hash_compact<<<(c->fw_htab + 255) / 256, 256, 0, st>>>(c->fw_ckeys, c->fw_cids, c->fw_htab, &c->fw_ctl->n_classes);
    LAUNCHED(c);
    // Steady state: the table buffer an earlier batch sized holds this batch's classes too.  Then nothing is read back
    // here: the class count stays on the device (grid sized for the buffer, surplus blocks leave), a class beyond the
    // buffer raises FAST_ERR_CAP, and the end of the evaluation (settle_locked) redoes such a batch with the counts read
    // back.  No host round trip in the middle of a batch, hence no idle GPU and, node-sharded, no skew between the shards.
    if (c->fw_opt_ok && c->fw_tables_classes > 0) {
        const u32 cap_now = std::min(ccap, c->fw_tables_classes);
        k_class_assign<<<pb, 256, 0, st>>>(c->fw_pods, P, c->d_max6, c->fw_crep, c->fw_cids, c->fw_slot, c->fw_cls_max, cap_now, c->fw_ctl);
        LAUNCHED(c);
        if (N) {
            k_class_order<<<(cap_now * 32 + CLS_ORDER_T - 1) / CLS_ORDER_T, CLS_ORDER_T, 0, st>>>(c->fw_cls_max, cap_now, c->fw_ctl, true, c->fw_cls_order);
            LAUNCHED(c);
            dim3 grid((N + CLST - 1) / CLST, (cap_now + CLS_PER - 1) / CLS_PER);
            k_class_tables<<<grid, CLST, 0, st>>>(c->snap, c->fw_cls_max, cap_now, c->fw_tables, c->fw_ctl, true, c->fw_cls_order);
            LAUNCHED(c);
        }
        c->fw_optimistic = true;
        if (c->fw_allow_defer)
            c->fw_deferred_err = true;
        else
            c->fw_gated = true;  // fw_opt_ok is only set with one of the two
        return YODA_OK;
    }
    k_class_assign<<<pb, 256, 0, st>>>(c->fw_pods, P, c->d_max6, c->fw_crep, c->fw_cids, c->fw_slot, c->fw_cls_max, ccap, c->fw_ctl);
    LAUNCHED(c);
    TRY(fast_read_ctl(c));
    c->fw_n_classes = c->fw_hctl->n_classes;
    if (c->fw_groups_pending) {
        c->fw_groups_pending = false;
        c->fw_n_groups = c->fw_hctl->n_groups;
        if (c->fw_n_groups > FAST_GROUP_CAP) {  // too many groups: the maxima of phase 0 are void, redo them exactly
            c->ph_path = 0;
            k_init_partials<<<(P + 255) / 256, 256, 0, st>>>(c->d_max6, c->d_count, c->d_hilo, c->d_ckey, P);
            LAUNCHED(c);
            TRY(direct_pods_launch(c, 0));
            // node-sharded: the maxima exchanged after phase 0 are void as well; every shard is here (the group count
            // depends on the pods alone), so they all repeat the exchange
            TRY(exchange_locked(c, 0));
            return YODA_OK;
        }
    }
    if (c->fw_hctl->err || c->fw_n_classes > ccap) {
        c->ph_path = 0;
        return YODA_OK;
    }
    if (c->fw_n_classes > c->fw_tables_classes) {
        graph_drop(c);
        dev_free(c->fw_tables);
        c->fw_tables_classes = 0;
        const u32 want = std::min(ccap, std::max(c->fw_n_classes * 2, 64u));
        TRY(dev_alloc(&c->fw_tables, (size_t)want * c->snap.cap * G));
        c->fw_tables_classes = want;
    }
    if (N) {
        k_class_order<<<(c->fw_n_classes * 32 + CLS_ORDER_T - 1) / CLS_ORDER_T, CLS_ORDER_T, 0, st>>>(c->fw_cls_max, c->fw_n_classes, c->fw_ctl, false, c->fw_cls_order);
        LAUNCHED(c);
        dim3 grid((N + CLST - 1) / CLST, (c->fw_n_classes + CLS_PER - 1) / CLS_PER);
        k_class_tables<<<grid, CLST, 0, st>>>(c->snap, c->fw_cls_max, c->fw_n_classes, c->fw_tables, c->fw_ctl, false, c->fw_cls_order);
        LAUNCHED(c);
    }
    if (c->fw_allow_defer) {
        c->fw_deferred_err = true;  // "every class fits the packed table": read back with the results
        return YODA_OK;
    }
    if (c->fw_gate_ok) {
        c->fw_gated = true;  // both paths are launched, the device picks (gate_closed); the verdict is read back with the results
        return YODA_OK;
    }
    TRY(fast_read_ctl(c));
    if (c->fw_hctl->err) {
        c->ph_path = 0;
        return YODA_OK;
    }
    return YODA_OK;
}

// ---- phases ---------------------------------------------------------------------------------
// phase 0: canonical pods, identities, local maxima (collection.CollectMaxValues)
static int phase0(yoda_ctx *c, const yoda_pod_req *d_req) {
    const u32 P = c->ph_P;
    cudaStream_t st = c->stream;
    c->launches = 0;
    TRY(ev_record(c, EV_START));
    c->fw_small = c->ph_path == 1 && P <= FAST_SMALL_P && P <= fast_class_cap(c);
    if (c->fw_small) {
        k_small_setup<<<1, 64, 0, st>>>(d_req, c->d_pods, P, c->d_max6, c->d_count, c->d_hilo, c->d_ckey, c->fw_gdesc, c->fw_pods, c->fw_m_sorted, c->fw_step,
                                        c->fw_ctl);
        LAUNCHED(c);
    } else {
        k_pod_prep<<<(P + 255) / 256, 256, 0, st>>>(d_req, c->d_pods, P);
        LAUNCHED(c);
        k_init_partials<<<(P + 255) / 256, 256, 0, st>>>(c->d_max6, c->d_count, c->d_hilo, c->d_ckey, P);
        LAUNCHED(c);
    }
    if (c->ph_path == 1) TRY(fast_phase0(c));
    if (c->ph_path == 1) TRY(wide_pods_launch(c, 0));  // after k_max_scan, which stores the table nodes' maxima
    if (c->ph_path == 0) {
        TRY(ev_record(c, EV_PREP));
        TRY(direct_pods_launch(c, 0));
    }
    TRY(ev_record(c, EV_MAX));
    return YODA_OK;
}

// phase 1: local lowest/highest over feasible nodes
static int phase1(yoda_ctx *c) {
    if (c->ph_path == 1) TRY(fast_classes(c));
    if (c->fw_agree) {
        // Sharded DECIDE: the one-pass variant (choices out of the score pass, one exchange less, no output pass)
        // packs the choice key differently from the exact kernels, so it is used only if every shard runs the
        // table path -- which a shard knows for itself only now.  One word, all-reduced with MAX: 1 = "I do not".
        cudaStream_t st = c->stream;
        CU(cudaMemsetAsync(c->d_agree, 0, sizeof(u64), st));
        if (c->ph_path != 1 || c->n_wide) CU(cudaMemsetAsync(c->d_agree, 1, 1, st));  // little-endian: the word becomes 1
        NC(g_nccl.AllReduce(c->d_agree, c->d_agree, 1, ncclUint64, ncclMax, c->comm, st));
        CU(cudaMemcpyAsync(c->h_agree, c->d_agree, sizeof(u64), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        c->fw_decide_merged = (*c->h_agree == 0);
    }
    if (c->ph_path == 0) c->fw_decide_merged = false;
    if (c->ph_path == 1) {
        TRY(fast_score_launch(c, 0, c->ph_P));
        TRY(wide_pods_launch(c, 1));
        if (c->fw_gated) TRY(direct_pods_launch(c, 1, true));
    }
    if (c->ph_path == 0) TRY(direct_pods_launch(c, 1));
    TRY(ev_record(c, EV_SCORE));
    return YODA_OK;
}

// phase 2: output pass (matrices when FULL), feasible counts, node choice
static int phase2(yoda_ctx *c) {
    if (c->ph_path == 1 && c->fw_decide_merged) return wide_pods_launch(c, 2);  // phase 1 already produced the table nodes' counts and choices
    if (c->ph_path == 1) {
        TRY(fast_out_launch(c, 0, c->ph_P));
        TRY(wide_pods_launch(c, 2));  // overwrites the wide nodes' matrix entries
        if (c->fw_gated) TRY(direct_out_launch(c, true));
        return YODA_OK;
    }
    return direct_out_launch(c);
}

static int phase3(yoda_ctx *c) {
    const u32 P = c->ph_P;
    k_finalize<<<(P + 255) / 256, 256, 0, c->stream>>>(c->d_max6, c->d_count, c->d_hilo, c->d_ckey, c->d_res, P,
                                                       c->ph_path == 1 && c->fw_decide_merged);
    LAUNCHED(c);
    TRY(ev_record(c, EV_WRITE));
    return YODA_OK;
}

static int begin_locked(yoda_ctx *c, u32 P, u32 flags) {
    if (c->flushed_uncommitted || c->n_upd)
        return fail(YODA_E_STATE, "uncommitted node updates are pending: call yoda_commit before evaluating");
    if (P > (1u << 22)) return fail(YODA_E_RANGE, "batch of %u pods exceeds the limit of %u per call; split it", P, 1u << 22);
    TRY(ensure_pods(c, P));
    c->stride = round_up(std::max(c->snap.n_nodes, 1u), 1024);  // rows padded to the hot kernels' CTA tile
    if (flags & ~(YODA_EVAL_FULL | YODA_EVAL_FORCE_DIRECT | YODA_EVAL_RAW32)) return fail(YODA_E_ARG, "unknown evaluation flags 0x%x", flags);
    c->ph_path = (c->table_ok && !(flags & YODA_EVAL_FORCE_DIRECT)) ? 1 : 0;
    // (a node-sharded FULL batch may pick the exact kernels on the device: room for int64 rows either way)
    if (flags & YODA_EVAL_FULL) TRY(ensure_matrices(c, P, c->stride, (c->ph_path == 1 && c->narrow_ok && (flags & YODA_EVAL_RAW32) && !c->comm) ? 4u : 8u));
    if (c->ph_path == 1) TRY(ensure_fast(c, P));
    c->ph_P = P;
    c->ph_flags = flags;
    c->fw_groups_pending = false;
    c->fw_gated = false;
    c->fw_gate_ok = false;
    c->fw_defer_groups = false;
    c->fw_opt_ok = false;
    c->fw_optimistic = false;
    c->g_replayed = false;
    c->ph_done = -1;
    c->last_P = 0;
    c->last_full = false;
    return YODA_OK;
}

// ---- exchanges over peer memory ---------------------------------------------------------------
// An all-reduce of one word over NCCL followed by a host wait: the barrier of the (rare) collective set-up steps.
static int nccl_host_barrier(yoda_ctx *c, u64 mine, int op_min, u64 *result) {
    CU(cudaMemcpyAsync(c->d_agree, &mine, sizeof(u64), cudaMemcpyHostToDevice, c->stream));
    NC(g_nccl.AllReduce(c->d_agree, c->d_agree, 1, ncclUint64, op_min ? ncclMin : ncclMax, c->comm, c->stream));
    CU(cudaMemcpyAsync(c->h_agree, c->d_agree, sizeof(u64), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (result) *result = *c->h_agree;
    return YODA_OK;
}

static void xchg_release(yoda_ctx *c) {
    for (u32 r = 0; r < X_MAX_RANKS; ++r) {
        if (c->x_opened[r]) cudaIpcCloseMemHandle(c->x_opened[r]);
        c->x_opened[r] = nullptr;
        c->x_peers.arena[r] = nullptr;
    }
    dev_free(c->x_arena);
    c->x_cap = 0;
    c->x_ok = false;
    cudaGetLastError();
}

// (Re)build the arenas for `cap` pods.  Collective: every shard calls it for the same batch.  Any failure on any shard
// (no IPC between these processes, no peer access, more than 8 shards), or YODA_B200_XCHG != peer, leaves ALL shards on NCCL.
static int xchg_setup(yoda_ctx *c, u32 cap) {
    const u32 R = c->cfg.shard_count, me = c->cfg.shard_rank;
    cudaStream_t st = c->stream;
    // nobody may still be reading an arena that is about to go away
    TRY(nccl_host_barrier(c, 0, 0, nullptr));
    xchg_release(c);
    // Opt-in (YODA_B200_XCHG=peer): measured on 8 B200s the one-shot exchange does not beat NCCL's all-reduce (config 3:
    // 8.38 vs 8.32 ms per step, config 4: 3.32 vs 3.31 ms) -- every shard pulls 7x the payload over NVLink where NCCL
    // reduces in two steps (and in the switch); it stays as the path without NCCL launches on the stream.
    const char *mode = getenv("YODA_B200_XCHG");
    u64 ok = (R <= X_MAX_RANKS && g_nccl.AllGather && mode && !strcmp(mode, "peer")) ? 1 : 0;
    cudaIpcMemHandle_t mine{}, all[X_MAX_RANKS]{};
    if (ok) {
        if (!c->x_handles && dev_alloc(&c->x_handles, (size_t)X_MAX_RANKS * sizeof(cudaIpcMemHandle_t))) ok = 0;
        if (ok && !c->x_done && (dev_alloc(&c->x_done, 1) || dev_alloc(&c->x_err, 1))) ok = 0;
        if (ok && dev_alloc(&c->x_arena, x_arena_words(cap))) ok = 0;
        if (ok && (cudaMemsetAsync(c->x_arena, 0, x_arena_words(cap) * sizeof(u64), st) != cudaSuccess ||
                   cudaMemsetAsync(c->x_done, 0, sizeof(u32), st) != cudaSuccess || cudaMemsetAsync(c->x_err, 0, sizeof(u32), st) != cudaSuccess ||
                   cudaIpcGetMemHandle(&mine, c->x_arena) != cudaSuccess))
            ok = 0;
        cudaGetLastError();
    }
    // every shard takes part in the gather whatever its own state (the collectives must match)
    if (g_nccl.AllGather && c->x_handles) {
        CU(cudaMemcpyAsync(c->x_handles + (size_t)me * sizeof mine, &mine, sizeof mine, cudaMemcpyHostToDevice, st));
        NC(g_nccl.AllGather(c->x_handles + (size_t)me * sizeof mine, c->x_handles, sizeof mine, ncclChar, c->comm, st));
        CU(cudaMemcpyAsync(all, c->x_handles, (size_t)R * sizeof mine, cudaMemcpyDeviceToHost, st));
    }
    u64 all_ok = 0;
    TRY(nccl_host_barrier(c, ok, 1, &all_ok));  // also: every arena is cleared before any peer raises a flag in it
    if (all_ok) {
        for (u32 r = 0; r < R && ok; ++r) {
            if (r == me) {
                c->x_peers.arena[r] = c->x_arena;
                continue;
            }
            void *p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                ok = 0;
                break;
            }
            c->x_opened[r] = p;
            c->x_peers.arena[r] = (u64 *)p;
        }
        TRY(nccl_host_barrier(c, ok, 1, &all_ok));
    }
    if (!all_ok) {
        xchg_release(c);
        return YODA_OK;  // NCCL carries the exchanges
    }
    c->x_cap = cap;
    c->x_ok = true;
    return YODA_OK;
}

// one partial buffer through the arenas: publish (the last buffer of the exchange raises the flags), then reduce
static int xchg_one(yoda_ctx *c, int phase, u64 *buf, size_t slot_off, size_t n, int op, bool last, bool reduce_now) {
    const u32 R = c->cfg.shard_count, me = c->cfg.shard_rank, par = (u32)(c->x_seq & 1u);
    const u64 tag = c->x_seq * 8 + c->x_gen[phase];
    const size_t voff = x_values_off(c->x_cap, par) + slot_off, foff = x_flags_off(c->x_cap, par, (u32)phase);
    const u32 blocks = (u32)std::min<size_t>((n + 255) / 256, 592);
    if (!reduce_now) {
        k_x_publish<<<blocks, 256, 0, c->stream>>>(buf, c->x_arena + voff, n, c->x_done, c->x_peers, foff, me, R, tag, last ? 1 : 0);
    } else {
        k_x_reduce<<<blocks, 256, 0, c->stream>>>(c->x_peers, voff, buf, n, op, c->x_arena + foff, R, tag, c->x_err);
    }
    LAUNCHED(c);
    return YODA_OK;
}

struct XBuf {
    u64 *buf;
    size_t off, n;
    int op;
};
static int xchg_phase(yoda_ctx *c, int phase, const XBuf *b, int nb) {
    for (int i = 0; i < nb; ++i) TRY(xchg_one(c, phase, b[i].buf, b[i].off, b[i].n, b[i].op, i == nb - 1, false));
    for (int i = 0; i < nb; ++i) TRY(xchg_one(c, phase, b[i].buf, b[i].off, b[i].n, b[i].op, false, true));
    c->x_gen[phase]++;
    return YODA_OK;
}

static int exchange_locked(yoda_ctx *c, int phase) {
    if (!c->comm) return YODA_OK;
    const u32 P = c->ph_P;
    if (c->x_ok) {
        // arena slots per parity, in pods: max6 [0, 6 cap), hilo [6 cap, 8 cap), ckey [8 cap, 9 cap), count [9 cap, 10 cap)
        const size_t cap = c->x_cap;
        const XBuf max6{c->d_max6, 0, (size_t)P * 6, 0}, hilo{c->d_hilo, cap * 6, (size_t)P * 2, 0}, ckey{c->d_ckey, cap * 8, P, 0},
            count{c->d_count, cap * 9, P, 1};
        if (phase == 0) {
            TRY(xchg_phase(c, 0, &max6, 1));
        } else if (phase == 1) {
            const XBuf all[3] = {hilo, ckey, count};
            TRY(xchg_phase(c, 1, all, c->fw_decide_merged ? 3 : 1));
        } else if (phase == 2 && !c->fw_decide_merged) {
            // with the counts left on the device, "a shard hit a capacity limit" rides along (settle_locked reads it)
            if (c->fw_optimistic) {
                k_cap_flag<<<1, 1, 0, c->stream>>>(c->fw_ctl, c->d_agree);
                LAUNCHED(c);
            }
            const XBuf all[3] = {ckey, count, XBuf{c->d_agree, cap * X_WORDS_PER_POD, 1, 0}};
            TRY(xchg_phase(c, 2, all, c->fw_optimistic ? 3 : 2));
        }
        return YODA_OK;
    }
    if (phase == 0) {
        NC(g_nccl.AllReduce(c->d_max6, c->d_max6, (size_t)P * 6, ncclUint64, ncclMax, c->comm, c->stream));
    } else if (phase == 1) {
        NC(g_nccl.GroupStart());
        NC(g_nccl.AllReduce(c->d_hilo, c->d_hilo, (size_t)P * 2, ncclUint64, ncclMax, c->comm, c->stream));
        if (c->fw_decide_merged) {  // agreed on by all shards in phase1: choice keys and counts are ready too
            NC(g_nccl.AllReduce(c->d_ckey, c->d_ckey, P, ncclUint64, ncclMax, c->comm, c->stream));
            NC(g_nccl.AllReduce(c->d_count, c->d_count, P, ncclUint64, ncclSum, c->comm, c->stream));
        }
        NC(g_nccl.GroupEnd());
    } else if (phase == 2 && !c->fw_decide_merged) {
        NC(g_nccl.GroupStart());
        NC(g_nccl.AllReduce(c->d_ckey, c->d_ckey, P, ncclUint64, ncclMax, c->comm, c->stream));
        NC(g_nccl.AllReduce(c->d_count, c->d_count, P, ncclUint64, ncclSum, c->comm, c->stream));
        NC(g_nccl.GroupEnd());
    }
    return YODA_OK;
}

static void collect_timing(yoda_ctx *c, bool copies) {
    yoda_timing &t = c->timing;
    memset(&t, 0, sizeof t);
    cudaEventElapsedTime(&t.total_ms, c->ev[EV_START], c->ev[EV_WRITE]);
    if (c->g_replayed) {  // one graph launch: the phases were not timed apart
        t.kernels = c->launches;
        t.path = (u32)c->ph_path;
        t.n_classes = c->fw_n_classes;
        t.n_groups = c->fw_n_groups;
        cudaGetLastError();
        return;
    }
    cudaEventElapsedTime(&t.prep_ms, c->ev[EV_START], c->ev[EV_PREP]);
    cudaEventElapsedTime(&t.max_ms, c->ev[EV_PREP], c->ev[EV_MAX]);
    cudaEventElapsedTime(&t.score_ms, c->ev[EV_MAX], c->ev[EV_SCORE]);
    cudaEventElapsedTime(&t.write_ms, c->ev[EV_SCORE], c->ev[EV_WRITE]);
    if (copies) {
        cudaEventElapsedTime(&t.h2d_ms, c->ev[EV_H2D0], c->ev[EV_H2D1]);
        cudaEventElapsedTime(&t.d2h_ms, c->ev[EV_D2H0], c->ev[EV_D2H1]);
    }
    t.kernels = c->launches;
    t.path = (u32)c->ph_path;
    t.n_classes = c->ph_path ? c->fw_n_classes : 0;
    t.n_groups = c->ph_path ? c->fw_n_groups : 0;
    cudaGetLastError();
}

static int run_all_locked(yoda_ctx *c, const yoda_pod_req *d_req) {
    if (c->comm) {
        // exchanges over peer memory: arenas sized like the pod buffers (every shard sees the same batches, so they all
        // grow at the same batch); a new batch flips the parity of the arena halves
        const char *xmode = getenv("YODA_B200_XCHG");
        if (!(xmode && !strcmp(xmode, "peer")) && c->x_cap == 0) c->x_cap = 0xFFFFFFFFu;  // NCCL (default): no set-up collectives at all
        if (c->ph_P > c->x_cap && (c->x_ok || c->x_cap == 0)) {
            TRY(xchg_setup(c, c->pods_cap));
            if (!c->x_ok) c->x_cap = 0xFFFFFFFFu;  // decided: NCCL; do not try again for every batch
        }
        c->x_seq++;
        c->x_gen[0] = c->x_gen[1] = c->x_gen[2] = 0;
    }
    c->fw_allow_defer = (c->comm == nullptr);
    c->fw_defer_groups = true;
    c->fw_gate_ok = (c->comm != nullptr) && (c->ph_flags & YODA_EVAL_FULL);
    c->fw_opt_ok = !c->fw_no_opt && (c->fw_allow_defer || c->fw_gate_ok);
    c->fw_optimistic = false;
    c->fw_deferred_err = false;
    c->fw_decide_merged = (c->comm == nullptr) && !(c->ph_flags & YODA_EVAL_FULL) && c->ph_path == 1;
    c->fw_agree = (c->comm != nullptr) && !(c->ph_flags & YODA_EVAL_FULL);
    TRY(phase0(c, d_req));
    TRY(exchange_locked(c, 0));
    TRY(phase1(c));
    TRY(exchange_locked(c, 1));
    TRY(phase2(c));
    TRY(exchange_locked(c, 2));
    TRY(phase3(c));
    if (c->comm && c->fw_optimistic && !c->x_ok) {  // every shard must take the same "redo" decision (settle_locked)
        k_cap_flag<<<1, 1, 0, c->stream>>>(c->fw_ctl, c->d_agree);
        LAUNCHED(c);
        NC(g_nccl.AllReduce(c->d_agree, c->d_agree, 1, ncclUint64, ncclMax, c->comm, c->stream));
    }
    c->fw_allow_defer = false;
    c->fw_defer_groups = false;
    c->fw_gate_ok = false;
    c->fw_opt_ok = false;
    c->fw_agree = false;
    c->ph_done = 3;
    c->last_P = c->ph_P;
    c->last_full = (c->ph_flags & YODA_EVAL_FULL) != 0;
    return YODA_OK;
}

// Small-batch path: the "every class fits the packed table" verdict was left on the device; `err` is
// its value.  If it failed, redo the batch with the exact kernels.
static int redo_direct_if_flagged(yoda_ctx *c, const yoda_pod_req *d_req, u32 err) {
    c->fw_deferred_err = false;
    if (!err) return YODA_OK;
    c->ph_path = 0;
    c->fw_decide_merged = false;
    TRY(phase0(c, d_req));
    TRY(phase1(c));
    TRY(phase2(c));
    TRY(phase3(c));
    return YODA_OK;
}

// A gated evaluation (see gate_closed) learns only now which kernels did the work: fw_hctl holds the verdict.
static void gate_verdict(yoda_ctx *c) {
    if (!c->fw_gated) return;
    c->fw_gated = false;
    if (c->fw_hctl->err) {  // the exact kernels: int64 rows
        c->ph_path = 0;
        c->res_raw32 = false;
    }
}

// After run_all_locked: bring the verdicts that were left on the device to the host (one synchronisation) and do what
// they ask for.  *redone = the evaluation was repeated (results must be fetched again).  Leaves the stream synchronised
// whenever it had anything to settle.
static int settle_locked(yoda_ctx *c, const yoda_pod_req *d_req, bool *redone) {
    *redone = false;
    for (int round = 0; round < 3; ++round) {
        if (!(c->fw_deferred_err || c->fw_gated || c->fw_optimistic)) {
            if (c->x_ok) {  // a shard that waited in vain for a peer must not hand out what it computed
                CU(cudaMemcpyAsync(&c->h_agree[1], c->x_err, sizeof(u32), cudaMemcpyDeviceToHost, c->stream));
                CU(cudaStreamSynchronize(c->stream));
                if ((u32)c->h_agree[1]) return fail(YODA_E_NCCL, "a shard did not arrive at a peer-memory exchange within 2 s");
            }
            return YODA_OK;
        }
        cudaStream_t st = c->stream;
        if (c->x_ok) CU(cudaMemcpyAsync(&c->h_agree[1], c->x_err, sizeof(u32), cudaMemcpyDeviceToHost, st));
        const bool collective = c->comm && c->fw_optimistic;
        CU(cudaMemcpyAsync(c->fw_hctl, c->fw_ctl, sizeof(FastCtl), cudaMemcpyDeviceToHost, st));
        if (collective) CU(cudaMemcpyAsync(c->h_agree, c->d_agree, sizeof(u64), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (c->x_ok && (u32)c->h_agree[1]) return fail(YODA_E_NCCL, "a shard did not arrive at a peer-memory exchange within 2 s");
        if (c->fw_optimistic) {  // the counts this evaluation ran without
            c->fw_n_classes = c->fw_hctl->n_classes;
            c->fw_n_groups = c->fw_hctl->n_groups;
            c->fw_groups_pending = false;
        }
        const bool over = c->fw_optimistic && ((c->fw_hctl->err & FAST_ERR_CAP) || (collective && *c->h_agree));
        c->fw_optimistic = false;
        if (over) {
            // more classes than the table buffer holds, or more groups than the limit (on some shard): once more with the
            // counts read back, which grows the buffer or hands the batch to the exact kernels
            c->fw_no_opt = true;
            c->ph_path = 1;
            c->fw_gated = false;
            c->fw_deferred_err = false;
            const int rc = run_all_locked(c, d_req);
            c->fw_no_opt = false;
            TRY(rc);
            *redone = true;
            continue;
        }
        gate_verdict(c);
        if (c->fw_deferred_err) {
            c->fw_deferred_err = false;
            if (c->fw_hctl->err) {
                TRY(redo_direct_if_flagged(c, d_req, c->fw_hctl->err));
                *redone = true;
            }
        }
        return YODA_OK;
    }
    return fail(YODA_E_STATE, "evaluation did not settle");
}

#define STAGE_PODS 4096u  // batches up to this size go through the library's pinned staging buffers
extern "C" int yoda_eval_batch(yoda_ctx *c, const yoda_pod_req *pods, u32 P, u32 flags, yoda_pod_result *out) {
    if (!c || (P && !pods)) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (P == 0) return YODA_OK;
    TRY(begin_locked(c, P, flags));
    cudaStream_t st = c->stream;
    const bool staged = P <= STAGE_PODS;  // small batches: one cheap host copy buys truly asynchronous DMA
    if (staged) memcpy(c->h_req, pods, (size_t)P * sizeof(yoda_pod_req));
    // ---- the one-pod scheduling cycle (and its <= 8-pod siblings) as a CUDA graph -----------------------------------
    if (c->g_enabled && !c->comm && c->ph_path == 1 && out && P <= FAST_SMALL_P && P <= c->fw_tables_classes && c->fw_ctl) {
        const yoda_ctx::GraphKey key{P, flags, c->snap.n_nodes, c->n_wide, st};
        if (!(c->g_exec && key == c->g_key) && key == c->g_seen) {
            // second cycle with this key (the first one ran call by call and left every buffer at its final size): capture
            if (c->g_exec) cudaGraphExecDestroy(c->g_exec);
            c->g_exec = nullptr;
            cudaGraph_t graph = nullptr;
            CU(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            c->g_capturing = true;
            int rc = YODA_OK;
            if (cudaMemcpyAsync(c->d_req, c->h_req, (size_t)P * sizeof(yoda_pod_req), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = YODA_E_CUDA;
            if (!rc) rc = run_all_locked(c, c->d_req);
            if (!rc && (cudaMemcpyAsync(c->h_res, c->d_res, (size_t)P * sizeof(yoda_pod_result), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                        cudaMemcpyAsync(c->fw_hctl, c->fw_ctl, sizeof(FastCtl), cudaMemcpyDeviceToHost, st) != cudaSuccess))  // results and verdict ride in the graph
                rc = YODA_E_CUDA;
            c->g_capturing = false;
            const cudaError_t ce = cudaStreamEndCapture(st, &graph);
            if (!rc && ce == cudaSuccess && graph && cudaGraphInstantiate(&c->g_exec, graph, 0) == cudaSuccess) {
                c->g_key = key;
                c->g_post = yoda_ctx::GraphPost{c->ph_path, c->ph_done, c->fw_small, c->fw_deferred_err, c->fw_decide_merged, c->res_raw32, c->last_full,
                                                c->last_P, c->fw_n_groups, c->fw_n_classes, c->launches};
            } else {
                c->g_exec = nullptr;
                c->g_enabled = false;  // this driver / sequence cannot be captured: call by call from now on
                cudaGetLastError();
                TRY(begin_locked(c, P, flags));
            }
            if (graph) cudaGraphDestroy(graph);
        }
        if (c->g_exec && key == c->g_key) {
            const yoda_ctx::GraphPost &gp = c->g_post;
            CU(cudaEventRecord(c->ev[EV_START], st));
            CU(cudaGraphLaunch(c->g_exec, st));
            CU(cudaEventRecord(c->ev[EV_WRITE], st));  // only the total is timed: copies included, no phase split
            c->g_replayed = true;
            c->ph_path = gp.ph_path; c->ph_done = gp.ph_done; c->fw_small = gp.fw_small; c->fw_deferred_err = gp.fw_deferred_err;
            c->fw_decide_merged = gp.fw_decide_merged; c->res_raw32 = gp.res_raw32; c->last_full = gp.last_full; c->last_P = gp.last_P;
            c->fw_n_groups = gp.fw_n_groups; c->fw_n_classes = gp.fw_n_classes; c->launches = gp.launches;
            CU(cudaStreamSynchronize(st));
            c->fw_deferred_err = false;
            if (c->fw_hctl->err) {  // the class tables do not fit the packed format: the cycle again with the exact kernels
                c->g_replayed = false;
                TRY(redo_direct_if_flagged(c, c->d_req, c->fw_hctl->err));
                CU(cudaMemcpyAsync(c->h_res, c->d_res, (size_t)P * sizeof(yoda_pod_result), cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
            }
            memcpy(out, c->h_res, (size_t)P * sizeof(yoda_pod_result));
            collect_timing(c, true);
            return YODA_OK;
        }
        c->g_seen = key;
    }
    CU(cudaEventRecord(c->ev[EV_H2D0], st));
    CU(cudaMemcpyAsync(c->d_req, staged ? c->h_req : pods, (size_t)P * sizeof(yoda_pod_req), cudaMemcpyHostToDevice, st));
    CU(cudaEventRecord(c->ev[EV_H2D1], st));
    TRY(run_all_locked(c, c->d_req));
    yoda_pod_result *dst = staged ? c->h_res : out;
    CU(cudaEventRecord(c->ev[EV_D2H0], st));
    if (out) CU(cudaMemcpyAsync(dst, c->d_res, (size_t)P * sizeof(yoda_pod_result), cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(c->ev[EV_D2H1], st));
    bool redone = false;
    TRY(settle_locked(c, c->d_req, &redone));  // the read-back of the verdicts rides on the same synchronisation as the results
    if (redone && out) CU(cudaMemcpyAsync(dst, c->d_res, (size_t)P * sizeof(yoda_pod_result), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (staged && out) memcpy(out, c->h_res, (size_t)P * sizeof(yoda_pod_result));
    collect_timing(c, true);
    return YODA_OK;
}

extern "C" int yoda_eval_batch_device(yoda_ctx *c, const void *d_pods, u32 P, u32 flags, void *d_out) {
    if (!c || (P && !d_pods)) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (P == 0) return YODA_OK;
    TRY(begin_locked(c, P, flags));
    TRY(run_all_locked(c, (const yoda_pod_req *)d_pods));
    bool redone = false;
    TRY(settle_locked(c, (const yoda_pod_req *)d_pods, &redone));
    if (d_out)
        CU(cudaMemcpyAsync(d_out, c->d_res, (size_t)P * sizeof(yoda_pod_result), cudaMemcpyDeviceToDevice, c->stream));
    return YODA_OK;
}

// ------------------------------------------------------------------------------------------
// sequential-commit batch (seq_kernels.cuh)
// ------------------------------------------------------------------------------------------
extern "C" int yoda_eval_sequential(yoda_ctx *c, const yoda_pod_req *pods, u32 P, const u32 *order, yoda_pod_result *out) {
    if (!c || (P && (!pods || !out))) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (P == 0) return YODA_OK;
    if (c->comm) return fail(YODA_E_STATE, "sequential-commit evaluation is not available on a node-sharded context");
    TRY(begin_locked(c, P, YODA_EVAL_DECIDE));
    if (c->ph_path != 1 || !c->narrow_ok)
        return fail(YODA_E_RANGE, "sequential-commit evaluation needs a narrow snapshot (card fields < 2^32-1, static scores < 2^24)");
    cudaStream_t st = c->stream;
    if (!c->sq_alloc) {
        TRY(dev_alloc(&c->sq_alloc, c->snap.cap));
        TRY(dev_alloc(&c->sq_stat, c->snap.cap));
        TRY(dev_alloc(&c->sq_status, 1));
    }
    if (P > c->sq_pods_cap) {
        CU(cudaStreamSynchronize(st));
        dev_free(c->sq_pos);
        dev_free(c->sq_order);
        c->sq_pods_cap = 0;
        TRY(dev_alloc(&c->sq_pos, c->pods_cap));
        TRY(dev_alloc(&c->sq_order, c->pods_cap));
        c->sq_pods_cap = c->pods_cap;
    }
    CU(cudaMemcpyAsync(c->d_req, pods, (size_t)P * sizeof(yoda_pod_req), cudaMemcpyHostToDevice, st));
    if (order) {
        for (u32 i = 0; i < P; ++i)
            if (order[i] >= P) return fail(YODA_E_ARG, "order[%u] = %u is not a pod of this batch", i, order[i]);
        CU(cudaMemcpyAsync(c->sq_order, order, (size_t)P * sizeof(u32), cudaMemcpyHostToDevice, st));
    }
    c->fw_allow_defer = false;
    c->fw_deferred_err = false;
    c->fw_decide_merged = true;  // the choice comes out as a packed (score, index) key, as in merged DECIDE
    // maxima, groups, classes and class tables do not depend on allocations: same kernels as the frozen path
    TRY(phase0(c, c->d_req));
    if (c->ph_path == 1) TRY(fast_classes(c));
    if (c->ph_path != 1)
        return fail(YODA_E_RANGE, "sequential-commit evaluation needs the table path (too many requirement groups or classes in this batch)");
    TRY(ev_record(c, EV_SCORE));
    const u32 N = c->snap.n_nodes;
    if (N) {
        k_seq_positions<<<(P + 255) / 256, 256, 0, st>>>(c->fw_pods, P, c->sq_pos);
        LAUNCHED(c);
        k_seq_init<<<(N + 255) / 256, 256, 0, st>>>(c->snap, c->sq_alloc, c->sq_stat);
        LAUNCHED(c);
        CU(cudaMemsetAsync(c->sq_status, 0, sizeof(u32), st));
        int per_sm = 0;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_seq_commit, SEQ_THREADS, 0));
        const u32 grid = std::max(1u, std::min<u32>((N + SEQ_THREADS - 1) / SEQ_THREADS, (u32)per_sm * (u32)c->n_sms));
        DevSnap snap = c->snap;
        const PodS *a_pods = c->fw_pods;
        const PodC *a_podc = c->d_pods;
        const u32 *a_pos = c->sq_pos, *a_order = order ? c->sq_order : nullptr, *a_tables = c->fw_tables;
        u32 a_P = P;
        void *args[] = {&snap, &a_pods, &a_podc, &a_pos, &a_order, &a_P, &a_tables, &c->sq_alloc, &c->sq_stat, &c->d_hilo, &c->d_count,
                        &c->d_ckey, &c->sq_status};
        CU(cudaLaunchCooperativeKernel((void *)k_seq_commit, dim3(grid), dim3(SEQ_THREADS), args, 0, st));
        LAUNCHED(c);
    }
    TRY(phase3(c));
    u32 status = 0;
    CU(cudaMemcpyAsync(out, c->d_res, (size_t)P * sizeof(yoda_pod_result), cudaMemcpyDeviceToHost, st));
    if (N) CU(cudaMemcpyAsync(&status, c->sq_status, sizeof(u32), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    c->fw_decide_merged = false;
    c->ph_done = 3;
    collect_timing(c, false);
    if (status) return fail(YODA_E_RANGE, "a committed allocation pushed a node's static score out of the narrow range");
    return YODA_OK;
}

extern "C" int yoda_sync(yoda_ctx *c) {
    if (!c) return fail(YODA_E_ARG, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    CU(cudaStreamSynchronize(c->stream));
    return YODA_OK;
}

extern "C" int yoda_last_timing(yoda_ctx *c, yoda_timing *t) {
    if (!c || !t) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (c->ph_done != 3) return fail(YODA_E_STATE, "no completed evaluation");
    CU(cudaStreamSynchronize(c->stream));
    const bool had_copies = c->timing.h2d_ms != 0.f;
    collect_timing(c, had_copies);
    *t = c->timing;
    return YODA_OK;
}

static int ensure_row_staging(yoda_ctx *c, size_t bytes) {
    if (bytes <= c->h_rows_cap) return YODA_OK;
    if (c->h_rows) cudaFreeHost(c->h_rows);
    c->h_rows = nullptr;
    c->h_rows_cap = 0;
    const size_t cap = std::max<size_t>(bytes, 1 << 16);
    if (cudaMallocHost((void **)&c->h_rows, cap) != cudaSuccess) {
        cudaGetLastError();
        return fail(YODA_E_NOMEM, "cudaMallocHost of %zu bytes of row staging failed", cap);
    }
    c->h_rows_cap = cap;
    return YODA_OK;
}

extern "C" int yoda_fetch_row(yoda_ctx *c, u32 pod, int kind, void *dst, size_t bytes) {
    if (!c || !dst) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (!c->last_full || pod >= c->last_P)
        return fail(YODA_E_STATE, "no YODA_EVAL_FULL result holds pod %u (last batch: %u pods, full=%d)", pod, c->last_P,
                    (int)c->last_full);
    const void *src;
    size_t row;
    switch (kind) {
        case YODA_ROW_FEASIBLE_BITS: src = c->d_bits + (size_t)pod * (c->stride / 32); row = (c->stride / 32) * 4; break;
        case YODA_ROW_RAW:
            if (c->res_raw32) {  // uint32 on the device, int64 for the caller: widened out of the pinned staging buffer
                if (bytes > c->stride * 8) return fail(YODA_E_RANGE, "row kind %d holds %zu bytes, %zu requested", kind, (size_t)c->stride * 8, bytes);
                const size_t n = bytes / 8;
                TRY(ensure_row_staging(c, n * 4));
                CU(cudaMemcpyAsync(c->h_rows, (const u32 *)c->d_raw + (size_t)pod * c->stride, n * 4, cudaMemcpyDeviceToHost, c->stream));
                CU(cudaStreamSynchronize(c->stream));
                const u32 *tmp = (const u32 *)c->h_rows;
                for (size_t i = 0; i < n; ++i) ((i64 *)dst)[i] = (i64)tmp[i];
                return YODA_OK;
            }
            src = c->d_raw + (size_t)pod * c->stride; row = c->stride * 8; break;
        case YODA_ROW_NORM: src = c->d_norm + (size_t)pod * c->stride; row = c->stride; break;
        default: return fail(YODA_E_ARG, "unknown row kind %d", kind);
    }
    if (bytes > row) return fail(YODA_E_RANGE, "row kind %d holds %zu bytes, %zu requested", kind, row, bytes);
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return YODA_OK;
}

// The three rows of one pod with a single synchronisation: three async copies into the pinned staging buffer
// (true DMA, whatever memory the caller's buffers are), then one pass that hands them out.
extern "C" int yoda_fetch_rows(yoda_ctx *c, u32 pod, u32 n_nodes, uint32_t *bits, int64_t *raw, uint8_t *norm) {
    if (!c || !bits || !raw || !norm) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (!c->last_full || pod >= c->last_P)
        return fail(YODA_E_STATE, "no YODA_EVAL_FULL result holds pod %u (last batch: %u pods, full=%d)", pod, c->last_P, (int)c->last_full);
    if (n_nodes > c->stride) return fail(YODA_E_RANGE, "%u nodes requested, rows hold %llu", n_nodes, (unsigned long long)c->stride);
    if (n_nodes == 0) return YODA_OK;
    const size_t words = (n_nodes + 31) / 32, raw_elem = c->res_raw32 ? 4 : 8;
    const size_t off_raw = (words * 4 + 15) & ~(size_t)15, off_norm = off_raw + (((size_t)n_nodes * raw_elem + 15) & ~(size_t)15);
    TRY(ensure_row_staging(c, off_norm + n_nodes));
    cudaStream_t st = c->stream;
    CU(cudaMemcpyAsync(c->h_rows, c->d_bits + (size_t)pod * (c->stride / 32), words * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(c->h_rows + off_raw, (const u8 *)c->d_raw + (size_t)pod * c->stride * raw_elem, (size_t)n_nodes * raw_elem, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(c->h_rows + off_norm, c->d_norm + (size_t)pod * c->stride, n_nodes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    memcpy(bits, c->h_rows, words * 4);
    if (c->res_raw32) {
        const u32 *src = (const u32 *)(c->h_rows + off_raw);
        for (u32 i = 0; i < n_nodes; ++i) raw[i] = (i64)src[i];
    } else {
        memcpy(raw, c->h_rows + off_raw, (size_t)n_nodes * 8);
    }
    memcpy(norm, c->h_rows + off_norm, n_nodes);
    return YODA_OK;
}

extern "C" int yoda_result_matrices(yoda_ctx *c, void **d_bits, void **d_raw, void **d_norm, uint64_t *stride) {
    if (!c) return fail(YODA_E_ARG, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->last_full) return fail(YODA_E_STATE, "no YODA_EVAL_FULL result");
    if (d_bits) *d_bits = c->d_bits;
    if (d_raw) *d_raw = c->d_raw;
    if (d_norm) *d_norm = c->d_norm;
    if (stride) *stride = c->stride;
    return YODA_OK;
}

// Introspection: the order in which the table path walked the pods of the last evaluation.
extern "C" int yoda_debug_pod_order(yoda_ctx *c, uint32_t *orig, uint32_t *group, uint32_t *cls, uint32_t n, uint32_t *score_chunk,
                                    uint32_t *out_chunk) {
    if (!c) return fail(YODA_E_ARG, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (c->ph_done != 3 || c->ph_path != 1 || n != c->ph_P) return fail(YODA_E_STATE, "the last evaluation did not take the table path with %u pods", n);
    std::vector<PodS> tmp(n);
    CU(cudaMemcpyAsync(tmp.data(), c->fw_pods, (size_t)n * sizeof(PodS), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    for (u32 i = 0; i < n; ++i) {
        if (orig) orig[i] = tmp[i].orig;
        if (group) group[i] = tmp[i].gid;
        if (cls) cls[i] = tmp[i].cls;
    }
    if (score_chunk) *score_chunk = c->last_score_chunk;
    if (out_chunk) *out_chunk = c->last_out_chunk;
    return YODA_OK;
}

extern "C" int yoda_result_raw_bytes(yoda_ctx *c, uint32_t *bytes) {
    if (!c || !bytes) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->last_full) return fail(YODA_E_STATE, "no YODA_EVAL_FULL result");
    *bytes = c->res_raw32 ? 4u : 8u;
    return YODA_OK;
}

// ------------------------------------------------------------------------------------------
// framework tail (tail_kernels.cuh): sampling, weight, seeded selectHost over the last FULL result
// ------------------------------------------------------------------------------------------
extern "C" uint32_t yoda_num_feasible_nodes_to_find(uint32_t num_all_nodes, int32_t percentage) {
    // [3P] k8s v1.20.0 generic_scheduler.go numFeasibleNodesToFind: minFeasibleNodesToFind 100, minimum percentage 5,
    // adaptive percentage 50 - nodes/125 when the configured one is <= 0 (deploy/yoda-scheduler.yaml:18 sets 0)
    const int32_t all = (int32_t)std::min<uint32_t>(num_all_nodes, 0x7FFFFFFFu);
    if (all < 100 || percentage >= 100) return num_all_nodes;
    int32_t adaptive = percentage;
    if (adaptive <= 0) adaptive = std::max(5, 50 - all / 125);
    const int32_t num = (int32_t)((int64_t)all * adaptive / 100);
    return num < 100 ? 100u : (uint32_t)num;
}

extern "C" int yoda_framework_tail(yoda_ctx *c, const yoda_tail_config *cfg, yoda_tail_result *out, uint32_t *next_start) {
    static_assert(sizeof(TailOut) == sizeof(yoda_tail_result), "TailOut mirrors yoda_tail_result");
    if (!c || !cfg || !out) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (!c->last_full || c->last_P == 0) return fail(YODA_E_STATE, "yoda_framework_tail works on the matrices of a YODA_EVAL_FULL batch: none is held");
    if (c->comm || c->cfg.shard_count != 1) return fail(YODA_E_STATE, "yoda_framework_tail is not available on a node-sharded context");
    if (cfg->flags & ~YODA_TAIL_CHAIN) return fail(YODA_E_ARG, "unknown tail flags 0x%x", cfg->flags);
    const u32 P = c->last_P, N = c->snap.n_nodes;
    cudaStream_t st = c->stream;
    if (N == 0) {
        for (u32 p = 0; p < P; ++p) {
            memset(&out[p], 0, sizeof out[p]);
            out[p].chosen = -1;
        }
        if (next_start) *next_start = 0;
        return YODA_OK;
    }
    const u32 tiles = (u32)(c->stride / TAIL_TILE);
    const u64 words = c->stride / 32;
    if ((size_t)P * tiles > c->tl_tilecnt_cap) {
        CU(cudaStreamSynchronize(st));
        dev_free(c->tl_tilecnt);
        c->tl_tilecnt_cap = 0;
        TRY(dev_alloc(&c->tl_tilecnt, (size_t)P * tiles));
        c->tl_tilecnt_cap = (size_t)P * tiles;
    }
    if (P > c->tl_out_cap) {
        CU(cudaStreamSynchronize(st));
        dev_free(c->tl_out);
        c->tl_out_cap = 0;
        TRY(dev_alloc(&c->tl_out, P));
        c->tl_out_cap = P;
    }
    const i64 *d_other = nullptr;
    if (cfg->other_scores) {
        if (N > c->tl_other_cap) {
            CU(cudaStreamSynchronize(st));
            dev_free(c->tl_other);
            c->tl_other_cap = 0;
            TRY(dev_alloc(&c->tl_other, c->snap.cap));
            c->tl_other_cap = c->snap.cap;
        }
        CU(cudaMemcpyAsync(c->tl_other, cfg->other_scores, (size_t)N * sizeof(i64), cudaMemcpyHostToDevice, st));
        d_other = c->tl_other;
    }
    const u32 want = yoda_num_feasible_nodes_to_find(N, cfg->percentage);
    const int chain = (cfg->flags & YODA_TAIL_CHAIN) ? 1 : 0;
    const u64 cells = (u64)P * tiles;
    k_tail_tilecnt<<<(unsigned)((cells + 255) / 256), 256, 0, st>>>(c->d_bits, words, P, tiles, c->tl_tilecnt);
    CU(cudaGetLastError());
    if (chain)
        k_tail_window<<<1, 32, 0, st>>>(c->d_bits, words, c->tl_tilecnt, tiles, P, N, want, cfg->start_index, 1, c->tl_out);
    else
        k_tail_window<<<(P + 7) / 8, 256, 0, st>>>(c->d_bits, words, c->tl_tilecnt, tiles, P, N, want, cfg->start_index, 0, c->tl_out);
    CU(cudaGetLastError());
    if (c->res_raw32)
        k_tail_pick<true><<<(P + 7) / 8, 256, 0, st>>>(c->d_bits, words, c->d_raw, c->stride, P, N, cfg->weight, cfg->seed, d_other, c->snap.node_base, c->tl_out);
    else
        k_tail_pick<false><<<(P + 7) / 8, 256, 0, st>>>(c->d_bits, words, c->d_raw, c->stride, P, N, cfg->weight, cfg->seed, d_other, c->snap.node_base, c->tl_out);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out, c->tl_out, (size_t)P * sizeof(TailOut), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (next_start) *next_start = chain ? (out[P - 1].start + out[P - 1].processed) % N : cfg->start_index % N;
    return YODA_OK;
}

// ------------------------------------------------------------------------------------------
// phase-split API (host-driven exchange)
// ------------------------------------------------------------------------------------------
extern "C" int yoda_phase_begin(yoda_ctx *c, const yoda_pod_req *pods, u32 P, u32 flags) {
    if (!c || !pods || P == 0) return fail(YODA_E_ARG, "null argument or empty batch");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    TRY(begin_locked(c, P, flags));
    c->fw_allow_defer = false;
    c->fw_deferred_err = false;
    c->fw_decide_merged = false;
    c->fw_agree = false;
    c->fw_groups_pending = false;
    CU(cudaMemcpyAsync(c->d_req, pods, (size_t)P * sizeof(yoda_pod_req), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));  // caller's buffer may be pageable and is not retained
    return YODA_OK;
}

extern "C" int yoda_phase_run(yoda_ctx *c, int phase) {
    if (!c) return fail(YODA_E_ARG, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (c->ph_P == 0 || phase != c->ph_done + 1 || phase > 3)
        return fail(YODA_E_STATE, "phase %d requested after phase %d", phase, c->ph_done);
    switch (phase) {
        case 0: TRY(phase0(c, c->d_req)); break;
        case 1: TRY(phase1(c)); break;
        case 2: TRY(phase2(c)); break;
        case 3: TRY(phase3(c)); break;
    }
    c->ph_done = phase;
    if (phase == 3) {
        c->last_P = c->ph_P;
        c->last_full = (c->ph_flags & YODA_EVAL_FULL) != 0;
    }
    CU(cudaStreamSynchronize(c->stream));  // exchange buffers are handed to a foreign collective next
    return YODA_OK;
}

extern "C" int yoda_phase_exchange(yoda_ctx *c, int phase, yoda_xchg *bufs, int *n_bufs) {
    if (!c || !bufs || !n_bufs) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    if (phase > c->ph_done) return fail(YODA_E_STATE, "phase %d has not run", phase);
    const u64 P = c->ph_P;
    int k = 0;
    auto put = [&](void *p, u64 n, int op) {
        if (k < *n_bufs) bufs[k] = yoda_xchg{p, n, op, 0};
        ++k;
    };
    if (phase == 0) {
        put(c->d_max6, P * 6, YODA_XCHG_MAX);
    } else if (phase == 1) {
        put(c->d_hilo, P * 2, YODA_XCHG_MAX);
    } else if (phase == 2) {
        put(c->d_ckey, P, YODA_XCHG_MAX);
        put(c->d_count, P, YODA_XCHG_SUM);
    }
    if (k > *n_bufs) return fail(YODA_E_RANGE, "need %d exchange descriptors", k);
    *n_bufs = k;
    return YODA_OK;
}

extern "C" int yoda_phase_finish(yoda_ctx *c, yoda_pod_result *out) {
    if (!c || !out) return fail(YODA_E_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (c->ph_done != 3) return fail(YODA_E_STATE, "phase 3 has not run");
    CU(cudaMemcpyAsync(out, c->d_res, (size_t)c->ph_P * sizeof(yoda_pod_result), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    collect_timing(c, false);
    return YODA_OK;
}

// ------------------------------------------------------------------------------------------
// NormalizeScore on a caller-supplied list
// ------------------------------------------------------------------------------------------
extern "C" int yoda_normalize_scores(yoda_ctx *c, int64_t *scores, u32 n) {
    if (!c || !scores) return fail(YODA_E_ARG, "null argument");
    if (n == 0) return fail(YODA_E_ARG, "empty score list (the reference indexes scores[0], scheduler.go:135)");
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    if (n > c->list_cap) {
        CU(cudaStreamSynchronize(c->stream));
        dev_free(c->d_list);
        c->list_cap = 0;
        TRY(dev_alloc(&c->d_list, round_up(n, 4096)));
        c->list_cap = round_up(n, 4096);
    }
    cudaStream_t st = c->stream;
    const i64 init[2] = {(i64)scores[0], 0};
    CU(cudaMemcpyAsync(c->d_list, scores, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(c->d_lohi, init, sizeof init, cudaMemcpyHostToDevice, st));
    const u32 blocks = std::min<u32>((n + 255) / 256, 1184);
    k_norm_list_reduce<<<blocks, 256, 0, st>>>(c->d_list, n, c->d_lohi);
    CU(cudaGetLastError());
    k_norm_list_apply<<<blocks, 256, 0, st>>>(c->d_list, n, c->d_lohi);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(scores, c->d_list, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return YODA_OK;
}

// ------------------------------------------------------------------------------------------
// QueueSort (sort.go:8-18): stable order of a pending batch by priority, higher first
// ------------------------------------------------------------------------------------------
__global__ void k_queue_keys(const i64 *__restrict__ pri, u32 n, u64 *keys, u32 *vals) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // descending priority as ascending unsigned key: flip the sign bit (signed -> unsigned order), then complement
    keys[i] = ~((u64)pri[i] ^ 0x8000000000000000ull);
    vals[i] = i;
}

extern "C" int yoda_queue_order(yoda_ctx *c, const int64_t *priority, u32 n, u32 *order) {
    if (!c || (n && (!priority || !order))) return fail(YODA_E_ARG, "null argument");
    if (n == 0) return YODA_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    cudaStream_t st = c->stream;
    u64 *keys_in = nullptr, *keys_out = nullptr;
    u32 *vals_in = nullptr, *vals_out = nullptr;
    i64 *d_pri = nullptr;
    void *tmp = nullptr;
    size_t bytes = 0;
    int rc = dev_alloc(&keys_in, n);
    if (!rc) rc = dev_alloc(&keys_out, n);
    if (!rc) rc = dev_alloc(&vals_in, n);
    if (!rc) rc = dev_alloc(&vals_out, n);
    if (!rc) rc = dev_alloc(&d_pri, n);
    auto done = [&](int code) {
        cudaStreamSynchronize(st);
        dev_free(keys_in); dev_free(keys_out); dev_free(vals_in); dev_free(vals_out); dev_free(d_pri);
        if (tmp) cudaFree(tmp);
        return code;
    };
    if (rc) return done(rc);
    if (cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, 64, st) != cudaSuccess ||
        cudaMalloc(&tmp, std::max<size_t>(bytes, 16)) != cudaSuccess)
        return done(fail(YODA_E_NOMEM, "sort scratch allocation failed"));
    if (cudaMemcpyAsync(d_pri, priority, (size_t)n * 8, cudaMemcpyHostToDevice, st) != cudaSuccess) return done(fail(YODA_E_CUDA, "H2D failed"));
    k_queue_keys<<<(n + 255) / 256, 256, 0, st>>>(d_pri, n, keys_in, vals_in);
    if (cub::DeviceRadixSort::SortPairs(tmp, bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, 64, st) != cudaSuccess)  // LSD radix: stable
        return done(fail(YODA_E_CUDA, "radix sort failed"));
    if (cudaMemcpyAsync(order, vals_out, (size_t)n * 4, cudaMemcpyDeviceToHost, st) != cudaSuccess) return done(fail(YODA_E_CUDA, "D2H failed"));
    if (cudaStreamSynchronize(st) != cudaSuccess) return done(fail(YODA_E_CUDA, "queue order failed: %s", cudaGetErrorString(cudaGetLastError())));
    return done(YODA_OK);
}

// ------------------------------------------------------------------------------------------
// label parsing (host): strconv.Atoi with error -> 0, then the uint/uint64 conversion
// ------------------------------------------------------------------------------------------
extern "C" uint64_t yoda_parse_label(const char *s, size_t len) {
    if (!s || len == 0) return 0;
    size_t i = 0;
    bool neg = false;
    if (s[0] == '+' || s[0] == '-') {
        neg = s[0] == '-';
        i = 1;
    }
    if (i == len) return 0;
    // accumulate in 128 bits so that overflow detection needs no pre-division
    unsigned __int128 acc = 0;
    const unsigned __int128 lim = neg ? ((unsigned __int128)1 << 63) : (((unsigned __int128)1 << 63) - 1);
    for (; i < len; ++i) {
        const unsigned char ch = (unsigned char)s[i];
        if (ch < '0' || ch > '9') return 0;  // syntax error
        acc = acc * 10 + (ch - '0');
        if (acc > lim) return 0;  // range error
    }
    const uint64_t mag = (uint64_t)acc;
    return neg ? (uint64_t)(0 - mag) : mag;
}

extern "C" int64_t yoda_uint64_to_int64(uint64_t v) { return (v >> 63) ? 0 : (int64_t)v; }

// ------------------------------------------------------------------------------------------
// NCCL communicator
// ------------------------------------------------------------------------------------------
extern "C" int yoda_comm_unique_id(void *id128) {
    if (!id128) return fail(YODA_E_ARG, "null argument");
    TRY(nccl_load());
    static_assert(sizeof(ncclUniqueId) == YODA_NCCL_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    NC(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return YODA_OK;
}

extern "C" int yoda_comm_init(yoda_ctx *c, const void *id128, u32 rank, u32 n_ranks) {
    if (!c || !id128) return fail(YODA_E_ARG, "null argument");
    if (n_ranks != c->cfg.shard_count || rank != c->cfg.shard_rank)
        return fail(YODA_E_ARG, "communicator rank %u/%u does not match the context's shard %u/%u", rank, n_ranks,
                    c->cfg.shard_rank, c->cfg.shard_count);
    TRY(nccl_load());
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard dg(c->cfg.device);
    ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    NC(g_nccl.CommInitRank(&c->comm, (int)n_ranks, id, (int)rank));
    if (!c->d_agree) {
        TRY(dev_alloc(&c->d_agree, 1));
        if (cudaMallocHost((void **)&c->h_agree, 2 * sizeof(u64)) != cudaSuccess) return fail(YODA_E_NOMEM, "cudaMallocHost of the agreement word failed");
        c->h_agree[0] = c->h_agree[1] = 0;  // [1]: error word of the peer-memory exchanges
    }
    return YODA_OK;
}

// ------------------------------------------------------------------------------------------
// SCV custom resource (JSON) -> yoda_node_record.  Host only.
// The CR's Go types live in github.com/NJUPT-ISL/SCV (un-vendored, go.mod:6); what the hot path reads of them is fixed
// by the reference's call sites: Status.{CardNumber, FreeMemorySum, TotalMemorySum, CardList[]} and
// Card.{Health, FreeMemory, TotalMemory, Clock, Core, Bandwidth, Power}.  The JSON spelling of those names cannot be
// checked offline, so keys are matched without regard to case ("cardList" == "CardList"); unknown keys are skipped.
// ------------------------------------------------------------------------------------------
namespace {
struct JsonCur {
    const char *p, *end;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
    bool eat(char ch) { ws(); if (p < end && *p == ch) { ++p; return true; } return false; }
};
static bool json_string(JsonCur &c, std::string *out) {
    if (!c.eat('"')) return false;
    std::string s;
    while (c.p < c.end && *c.p != '"') {
        if (*c.p == '\\') {
            if (++c.p >= c.end) return false;
            switch (*c.p) {
                case 'n': s += '\n'; break; case 't': s += '\t'; break; case 'r': s += '\r'; break; case 'b': s += '\b'; break; case 'f': s += '\f'; break;
                case 'u': if (c.end - c.p < 5) return false; s += '?'; c.p += 4; break;  // not needed for the fields read here
                default: s += *c.p;
            }
            ++c.p;
        } else {
            s += *c.p++;
        }
    }
    if (c.p >= c.end) return false;
    ++c.p;
    if (out) *out = std::move(s);
    return true;
}
static bool json_skip(JsonCur &c, int depth = 0);
static bool json_uint(JsonCur &c, uint64_t *out) {  // a JSON number without sign, fraction or exponent (what encoding/json writes for uint)
    c.ws();
    if (c.p >= c.end || *c.p < '0' || *c.p > '9') return false;
    unsigned __int128 v = 0;
    while (c.p < c.end && *c.p >= '0' && *c.p <= '9') {
        v = v * 10 + (unsigned)(*c.p++ - '0');
        if (v > 0xFFFFFFFFFFFFFFFFull) return false;
    }
    if (c.p < c.end && (*c.p == '.' || *c.p == 'e' || *c.p == 'E')) return false;
    *out = (uint64_t)v;
    return true;
}
static bool json_skip(JsonCur &c, int depth) {
    if (depth > 64) return false;
    c.ws();
    if (c.p >= c.end) return false;
    if (*c.p == '"') return json_string(c, nullptr);
    if (*c.p == '{' || *c.p == '[') {
        const char close = *c.p == '{' ? '}' : ']';
        const bool obj = *c.p == '{';
        ++c.p;
        if (c.eat(close)) return true;
        do {
            if (obj && (!json_string(c, nullptr) || !c.eat(':'))) return false;
            if (!json_skip(c, depth + 1)) return false;
        } while (c.eat(','));
        return c.eat(close);
    }
    while (c.p < c.end && *c.p != ',' && *c.p != '}' && *c.p != ']' && *c.p != ' ' && *c.p != '\n' && *c.p != '\t' && *c.p != '\r') ++c.p;  // number, true, false, null
    return true;
}
static bool key_is(const std::string &k, const char *name) {
    size_t i = 0;
    for (; i < k.size() && name[i]; ++i)
        if (tolower((unsigned char)k[i]) != tolower((unsigned char)name[i])) return false;
    return i == k.size() && !name[i];
}
// visits "key": value pairs of the object at the cursor; f returns false on a malformed value
template <typename F>
static bool json_object(JsonCur &c, F f) {
    if (!c.eat('{')) return false;
    if (c.eat('}')) return true;
    do {
        std::string k;
        if (!json_string(c, &k) || !c.eat(':')) return false;
        if (!f(k)) return false;
    } while (c.eat(','));
    return c.eat('}');
}
}  // namespace

extern "C" int yoda_record_from_scv_json(const char *json, size_t len, yoda_node_record *rec, char *name, size_t name_cap) {
    if (!json || !rec) return fail(YODA_E_ARG, "null argument");
    memset(rec, 0, sizeof *rec);
    if (name && name_cap) name[0] = 0;
    bool too_many = false, saw_status_fields = false;
    auto card = [&](JsonCur &c, uint32_t g) {
        return json_object(c, [&](const std::string &k) {
            uint64_t v = 0;
            if (key_is(k, "health")) {
                std::string h;
                if (!json_string(c, &h)) return false;
                if (g < YODA_MAX_CARDS && h == "Healthy") rec->healthy_mask |= 1u << g;  // filter.go:53,57: exactly this string
                return true;
            }
            uint64_t *dst = nullptr;
            if (g < YODA_MAX_CARDS) {
                if (key_is(k, "freeMemory")) dst = &rec->free_memory[g];
                else if (key_is(k, "totalMemory")) dst = &rec->total_memory[g];
                else if (key_is(k, "clock")) dst = &rec->clock[g];
                else if (key_is(k, "core")) dst = &rec->core[g];
                else if (key_is(k, "bandwidth")) dst = &rec->bandwidth[g];
                else if (key_is(k, "power")) dst = &rec->power[g];
            }
            if (!dst) return json_skip(c);
            if (!json_uint(c, &v)) return false;
            *dst = v;
            return true;
        });
    };
    std::function<bool(JsonCur &)> status = [&](JsonCur &c) {
        return json_object(c, [&](const std::string &k) {
            uint64_t v = 0;
            if (key_is(k, "cardList")) {
                saw_status_fields = true;
                if (c.eat('n')) { c.p += 3; return true; }  // null: no cards
                if (!c.eat('[')) return false;
                uint32_t g = 0;
                if (!c.eat(']')) {
                    do {
                        if (!card(c, g)) return false;
                        ++g;
                    } while (c.eat(','));
                    if (!c.eat(']')) return false;
                }
                too_many = g > YODA_MAX_CARDS;
                rec->n_cards = g;
                return true;
            }
            uint64_t *dst = key_is(k, "cardNumber") ? &rec->card_number : key_is(k, "freeMemorySum") ? &rec->free_memory_sum
                            : key_is(k, "totalMemorySum") ? &rec->total_memory_sum : nullptr;
            if (!dst) return json_skip(c);
            saw_status_fields = true;
            if (!json_uint(c, &v)) return false;
            *dst = v;
            return true;
        });
    };
    JsonCur c{json, json + len};
    // either the whole object ({"metadata": {"name": ...}, "status": {...}}) or the status alone
    const char *start = c.p;
    bool has_status_key = false;
    bool ok = json_object(c, [&](const std::string &k) {
        if (key_is(k, "status")) {
            has_status_key = true;
            return status(c);
        }
        if (key_is(k, "metadata"))
            return json_object(c, [&](const std::string &mk) {
                if (!key_is(mk, "name")) return json_skip(c);
                std::string n;
                if (!json_string(c, &n)) return false;
                if (name && name_cap) snprintf(name, name_cap, "%s", n.c_str());
                return true;
            });
        return json_skip(c);
    });
    if (ok && !has_status_key) {
        JsonCur d{start, json + len};
        memset(rec, 0, sizeof *rec);
        ok = status(d);
    }
    if (!ok) return fail(YODA_E_ARG, "malformed Scv JSON near offset %zu", (size_t)(c.p - json));
    if (!saw_status_fields) return fail(YODA_E_ARG, "no Scv status fields (cardList, cardNumber, freeMemorySum, totalMemorySum) found");
    if (too_many) return fail(YODA_E_RANGE, "the Scv lists %u cards, the device snapshot holds at most %d per node", rec->n_cards, YODA_MAX_CARDS);
    return YODA_OK;
}
