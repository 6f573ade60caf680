// framework.hpp -- the slice of the Kubernetes scheduling framework (k8s.io/kubernetes v1.20
// pkg/scheduler/framework) and of the SCV API (github.com/NJUPT-ISL/SCV api/v1) that the Yoda
// plugin touches on its Filter/Score path, restated as plain C++ types so that the plugin mirror
// and the benchmark harness can run without a cluster ("fake framework handle").
// Only what pkg/yoda/scheduler.go uses is here; names follow the Go originals.
#pragma once
#include <atomic>
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace v1 {
struct Pod {
    std::string Name, Namespace;
    std::map<std::string, std::string> Labels;
    const std::map<std::string, std::string> &GetLabels() const { return Labels; }
};
struct Node {
    std::string Name;
};
}  // namespace v1

namespace scv {  // [3P] SCV api/v1: fields as forced by the reference's call sites
struct Card {
    std::string Health;
    uint64_t FreeMemory = 0, TotalMemory = 0;
    uint64_t Clock = 0, Core = 0, Bandwidth = 0, Power = 0;  // Go uint (64-bit on amd64)
};
struct ScvStatus {
    uint64_t CardNumber = 0;
    std::vector<Card> CardList;
    uint64_t FreeMemorySum = 0, TotalMemorySum = 0;
};
struct Scv {
    std::string Name;  // cluster-scoped, named after the node (scheduler.go:80)
    ScvStatus Status;
};
}  // namespace scv

namespace framework {
constexpr int64_t MaxNodeScore = 100;
enum Code { Success = 0, Error = 1, Unschedulable = 2 };
struct Status {
    Code code = Success;
    std::string message;
    bool IsSuccess() const { return code == Success; }
};
inline Status NewStatus(Code c, std::string msg = "") { return Status{c, std::move(msg)}; }

struct StateData {
    virtual ~StateData() = default;
};
// framework.CycleState: per-scheduling-cycle key/value store with its own read/write lock (Go: sync.RWMutex)
class CycleState {
   public:
    CycleState() : serial_(next_serial().fetch_add(1, std::memory_order_relaxed) + 1) {}
    // Identity of this cycle: unique per CycleState object for the life of the process (an address would be reused).
    // Not part of the Go type; the C++ mirror uses it to remember, per thread, what it already looked up in this cycle.
    uint64_t Serial() const { return serial_; }
    void Lock() { mu_.lock(); }
    void Unlock() { mu_.unlock(); }
    void RLock() { mu_.lock_shared(); }
    void RUnlock() { mu_.unlock_shared(); }
    void Write(const std::string &key, std::shared_ptr<StateData> v) { data_[key] = std::move(v); }
    std::shared_ptr<StateData> Read(const std::string &key) const {  // nullptr == "not found" error
        auto it = data_.find(key);
        return it == data_.end() ? nullptr : it->second;
    }
    // the same without taking a reference: the entry lives as long as the cycle state does
    const StateData *Peek(const std::string &key) const {
        auto it = data_.find(key);
        return it == data_.end() ? nullptr : it->second.get();
    }

   private:
    static std::atomic<uint64_t> &next_serial() {
        static std::atomic<uint64_t> n{0};
        return n;
    }
    const uint64_t serial_;
    std::shared_mutex mu_;
    std::unordered_map<std::string, std::shared_ptr<StateData>> data_;
};

struct PodInfo {
    v1::Pod Pod;
};
struct NodeInfo {
    v1::Node node;
    std::vector<PodInfo> Pods;  // pods running on the node (algorithm.go:76)
    const v1::Node &Node() const { return node; }
};
struct QueuedPodInfo {
    v1::Pod Pod;
};
struct NodeScore {
    std::string Name;
    int64_t Score = 0;
};
using NodeScoreList = std::vector<NodeScore>;
struct PostFilterResult {};

// framework.Handle, reduced to SnapshotSharedLister().NodeInfos().Get (scheduler.go:111)
class Handle {
   public:
    virtual ~Handle() = default;
    virtual const NodeInfo *GetNodeInfo(const std::string &nodeName) const = 0;
};
}  // namespace framework
