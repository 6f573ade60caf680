// yoda_plugin.hpp -- C++ mirror of the reference's plugin surface (pkg/yoda/scheduler.go) on top of
// the C ABI of include/yoda_b200.h.  Same names, argument meaning and error behaviour as the Go
// plugin; the bodies talk to the device snapshot instead of a controller-runtime cache.
// This is the host layer the Go plugin would be (the Go toolchain is absent from the build image;
// the equivalent Go source is under go/).
#pragma once
#include <optional>
#include <shared_mutex>

#include "../../include/yoda_b200.h"
#include "framework.hpp"

namespace yoda {
constexpr const char *Name = "yoda";  // scheduler.go:23-25

// collection.Data / collection.MaxValue (collection.go:10-21), stored under CycleState key "Max"
struct MaxValue {
    uint64_t MaxBandwidth = 1, MaxClock = 1, MaxCore = 1, MaxFreeMemory = 1, MaxPower = 1, MaxTotalMemory = 1;
};
struct Data : framework::StateData {
    MaxValue Value;
};

// strconv.Atoi: value and error kind (0 ok, 1 syntax, 2 range; range errors return the clamped value)
std::pair<int64_t, int> Atoi(const std::string &s);
uint64_t StrToUint64(const std::string &s);  // filter.go:60-74
int64_t Uint64ToInt64(uint64_t v);           // filter.go:84-86

class Yoda {
   public:
    struct Options {
        int device = 0;
        uint32_t max_nodes = 1 << 16;
    };
    // scheduler.go:46-74 New(_ runtime.Object, h framework.Handle); error -> nullptr + message
    static std::unique_ptr<Yoda> New(const Options &opt, const framework::Handle *h, std::string *err = nullptr);
    ~Yoda();

    std::string Name() const { return yoda::Name; }  // scheduler.go:42-44

    // ---- what the informers deliver (replaces the controller-runtime cache, scheduler.go:53-73) ----
    framework::Status OnScvUpsert(const scv::Scv &s);
    void OnScvDelete(const std::string &nodeName);
    void OnPodAdd(const std::string &nodeName, const v1::Pod &pod);     // pod bound to / running on the node
    void OnPodDelete(const std::string &nodeName, const v1::Pod &pod);
    framework::Status Sync();  // commit staged updates to the device snapshot

    // ---- plugin methods, signatures as scheduler.go:76-161 ----
    framework::Status Filter(framework::CycleState &state, const v1::Pod &pod, const framework::NodeInfo &node);
    std::pair<framework::PostFilterResult, framework::Status> PostFilter(framework::CycleState &state, const v1::Pod &pod);
    bool Less(const framework::QueuedPodInfo &a, const framework::QueuedPodInfo &b) const;
    std::pair<int64_t, framework::Status> Score(framework::CycleState &state, const v1::Pod &p, const std::string &nodeName);
    framework::Status NormalizeScore(framework::CycleState &state, const v1::Pod &pod, framework::NodeScoreList &scores);
    Yoda *ScoreExtensions() { return this; }

    // ---- batch evaluation used by the benchmark harness: one device pass for all pending pods ----
    framework::Status EvalBatch(const std::vector<v1::Pod> &pods, bool full, std::vector<yoda_pod_result> *out);
    // one pod after the other in sort.Less order, every choice assumed onto its node before the next pod is scored
    framework::Status EvalSequential(const std::vector<v1::Pod> &pods, std::vector<yoda_pod_result> *out, std::vector<uint32_t> *order = nullptr);
    int FetchRow(uint32_t pod, int kind, void *dst, size_t bytes);
    static yoda_pod_req Requirement(const v1::Pod &pod);  // label parsing, filter.go:11-50
    int32_t NodeIndex(const std::string &name) const;     // -1 if never seen
    std::string NodeName(uint32_t idx) const {
        std::shared_lock<std::shared_mutex> lk(mu_);
        return names_[idx];
    }
    uint32_t NodeCount() const {
        std::shared_lock<std::shared_mutex> lk(mu_);
        return (uint32_t)names_.size();
    }
    yoda_ctx *ctx() { return ctx_; }

   private:
    Yoda() = default;
    struct Memo;  // per-cycle memo kept in CycleState
    const Memo *memo(framework::CycleState &state, const v1::Pod &pod, framework::Status *st);
    uint32_t index_of(const std::string &name);
    using IndexMap = std::unordered_map<std::string, uint32_t>;

    const framework::Handle *handle_ = nullptr;
    yoda_ctx *ctx_ = nullptr;
    // mu_ guards everything below: event handlers and evaluations take it exclusively.  The per-node calls of a cycle
    // (Filter / Score from 16 framework threads at once) do not take it at all: they read the cycle's memo, which
    // pins the name -> index map as it was at the evaluation.  The map is copied before a new node is added to it
    // while some memo still pins it, and changed in place otherwise.
    mutable std::shared_mutex mu_;
    std::shared_ptr<IndexMap> index_ = std::make_shared<IndexMap>();
    std::vector<std::string> names_;
    std::vector<uint8_t> has_scv_;
    std::vector<std::string> reject_;  // why the node's last Scv was refused ("" = accepted)
    uint32_t n_rejected_ = 0;          // non-empty entries of reject_
    std::vector<uint64_t> alloc_;
    bool dirty_ = false;
    framework::Status sync_locked();
};

// sort.Less / sort.GetPodPriority (sort.go:8-18)
int64_t GetPodPriority(const framework::QueuedPodInfo &p);
}  // namespace yoda
