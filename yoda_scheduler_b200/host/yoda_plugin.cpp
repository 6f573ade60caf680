// yoda_plugin.cpp -- see yoda_plugin.hpp.  Citations: pkg/yoda/scheduler.go of the reference.
#include "yoda_plugin.hpp"

#include <algorithm>
#include <cstring>

namespace yoda {
using framework::NewStatus;
using framework::Status;

// ---- labels (filter.go:60-86, sort.go:12-18) -------------------------------------------------
std::pair<int64_t, int> Atoi(const std::string &s) {
    if (s.empty()) return {0, 1};
    size_t i = 0;
    bool neg = false;
    if (s[0] == '+' || s[0] == '-') {
        neg = s[0] == '-';
        i = 1;
    }
    if (i == s.size()) return {0, 1};
    unsigned __int128 acc = 0;
    bool range = false;
    for (; i < s.size(); ++i) {
        const unsigned char ch = (unsigned char)s[i];
        if (ch < '0' || ch > '9') return {0, 1};
        if (!range) {
            acc = acc * 10 + (ch - '0');
            if (acc > ((unsigned __int128)1 << 63)) range = true;
        }
    }
    if (!neg && acc > (((unsigned __int128)1 << 63) - 1)) range = true;
    if (range) return {neg ? INT64_MIN : INT64_MAX, 2};
    const uint64_t mag = (uint64_t)acc;
    return {neg ? (int64_t)(0 - mag) : (int64_t)mag, 0};
}
uint64_t StrToUint64(const std::string &s) { return yoda_parse_label(s.data(), s.size()); }
int64_t Uint64ToInt64(uint64_t v) { return yoda_uint64_to_int64(v); }

int64_t GetPodPriority(const framework::QueuedPodInfo &p) {
    auto it = p.Pod.Labels.find("scv/priority");
    return it == p.Pod.Labels.end() ? 0 : Atoi(it->second).first;  // the error is dropped (sort.go:14)
}

yoda_pod_req Yoda::Requirement(const v1::Pod &pod) {
    yoda_pod_req r{};
    const auto &l = pod.GetLabels();
    auto it = l.find("scv/number");
    if (it != l.end()) { r.flags |= YODA_POD_HAS_NUMBER; r.number = StrToUint64(it->second); }
    it = l.find("scv/memory");
    if (it != l.end()) { r.flags |= YODA_POD_HAS_MEMORY; r.memory = StrToUint64(it->second); }
    it = l.find("scv/clock");
    if (it != l.end()) { r.flags |= YODA_POD_HAS_CLOCK; r.clock = StrToUint64(it->second); }
    return r;
}

// ---- construction ----------------------------------------------------------------------------
std::unique_ptr<Yoda> Yoda::New(const Options &opt, const framework::Handle *h, std::string *err) {
    std::unique_ptr<Yoda> y(new Yoda());
    y->handle_ = h;
    yoda_config cfg{};
    cfg.device = opt.device;
    cfg.max_nodes = opt.max_nodes;
    cfg.shard_count = 1;
    if (yoda_create(&cfg, &y->ctx_) != YODA_OK) {
        if (err) *err = yoda_last_error();
        return nullptr;
    }
    return y;
}
Yoda::~Yoda() { yoda_destroy(ctx_); }

uint32_t Yoda::index_of(const std::string &name) {
    auto it = index_->find(name);
    if (it != index_->end()) return it->second;
    const uint32_t idx = (uint32_t)names_.size();
    if (index_.use_count() > 1) index_ = std::make_shared<IndexMap>(*index_);  // a cycle's memo still reads the old one
    index_->emplace(name, idx);
    names_.push_back(name);
    has_scv_.push_back(0);
    reject_.emplace_back();
    alloc_.push_back(0);
    return idx;
}
int32_t Yoda::NodeIndex(const std::string &name) const {
    std::shared_lock<std::shared_mutex> lk(mu_);
    auto it = index_->find(name);
    return it == index_->end() ? -1 : (int32_t)it->second;
}

// ---- informer events ---------------------------------------------------------------------------
Status Yoda::OnScvUpsert(const scv::Scv &s) {
    std::unique_lock<std::shared_mutex> lk(mu_);
    const uint32_t idx = index_of(s.Name);
    yoda_node_record rec{};
    rec.card_number = s.Status.CardNumber;
    rec.free_memory_sum = s.Status.FreeMemorySum;
    rec.total_memory_sum = s.Status.TotalMemorySum;
    rec.n_cards = (uint32_t)s.Status.CardList.size();
    // A refused object (more than 8 cards; TotalMemorySum == 0, where the reference would panic) must not leave the
    // node answering from its previous record: it loses its SCV, with the reason kept for Filter's message.
    auto refuse = [&](const std::string &why) {
        yoda_node_remove(ctx_, idx);
        has_scv_[idx] = 0;
        n_rejected_ += reject_[idx].empty() ? 1 : 0;
        reject_[idx] = why;
        dirty_ = true;
        return NewStatus(framework::Error, why);
    };
    if (rec.n_cards > YODA_MAX_CARDS) return refuse("Node:" + s.Name + " has more than 8 cards");
    for (uint32_t g = 0; g < rec.n_cards; ++g) {
        const scv::Card &c = s.Status.CardList[g];
        if (c.Health == "Healthy") rec.healthy_mask |= 1u << g;  // filter.go:53,57
        rec.free_memory[g] = c.FreeMemory;
        rec.total_memory[g] = c.TotalMemory;
        rec.clock[g] = c.Clock;
        rec.core[g] = c.Core;
        rec.bandwidth[g] = c.Bandwidth;
        rec.power[g] = c.Power;
    }
    if (yoda_node_upsert(ctx_, idx, &rec) != YODA_OK) return refuse(yoda_last_error());
    has_scv_[idx] = 1;
    n_rejected_ -= reject_[idx].empty() ? 0 : 1;
    reject_[idx].clear();
    dirty_ = true;
    return NewStatus(framework::Success);
}
void Yoda::OnScvDelete(const std::string &nodeName) {
    std::unique_lock<std::shared_mutex> lk(mu_);
    const uint32_t idx = index_of(nodeName);
    yoda_node_remove(ctx_, idx);
    has_scv_[idx] = 0;
    n_rejected_ -= reject_[idx].empty() ? 0 : 1;
    reject_[idx].clear();
    dirty_ = true;
}
static uint64_t pod_memory(const v1::Pod &pod, bool *has) {
    auto it = pod.Labels.find("scv/memory");  // algorithm.go:77
    *has = it != pod.Labels.end();
    return *has ? StrToUint64(it->second) : 0;
}
void Yoda::OnPodAdd(const std::string &nodeName, const v1::Pod &pod) {
    bool has;
    const uint64_t m = pod_memory(pod, &has);
    if (!has) return;
    std::unique_lock<std::shared_mutex> lk(mu_);
    const uint32_t idx = index_of(nodeName);
    alloc_[idx] += m;  // wraps like the Go uint64 sum (algorithm.go:78)
    yoda_node_set_alloc(ctx_, idx, alloc_[idx]);
    dirty_ = true;
}
void Yoda::OnPodDelete(const std::string &nodeName, const v1::Pod &pod) {
    bool has;
    const uint64_t m = pod_memory(pod, &has);
    if (!has) return;
    std::unique_lock<std::shared_mutex> lk(mu_);
    const uint32_t idx = index_of(nodeName);
    alloc_[idx] -= m;
    yoda_node_set_alloc(ctx_, idx, alloc_[idx]);
    dirty_ = true;
}
Status Yoda::Sync() {
    std::unique_lock<std::shared_mutex> lk(mu_);
    return sync_locked();
}
Status Yoda::sync_locked() {
    if (dirty_) {
        if (yoda_commit(ctx_, nullptr) != YODA_OK) return NewStatus(framework::Error, yoda_last_error());
        dirty_ = false;
    }
    return NewStatus(framework::Success);
}

// ---- per-cycle memo ----------------------------------------------------------------------------
struct Yoda::Memo : framework::StateData {
    yoda_pod_result res{};
    uint32_t n_nodes = 0;            // nodes the evaluation covered; indices >= n_nodes were registered later
    std::vector<uint8_t> has;        // has_scv_[0..n_nodes) as of the evaluation: the cycle's frozen view
    std::shared_ptr<const IndexMap> index;  // name -> index as of the evaluation (never modified while pinned here)
    int32_t index_of(const std::string &name) const {
        auto it = index->find(name);
        return it == index->end() ? -1 : (int32_t)it->second;
    }
    std::unordered_map<uint32_t, std::string> reject;  // nodes whose last Scv was refused, and why (usually none)
    const std::string *rejected(uint32_t idx) const {
        auto it = reject.find(idx);
        return it == reject.end() ? nullptr : &it->second;
    }
    std::vector<uint32_t> bits;
    std::vector<int64_t> raw;
    std::vector<uint8_t> norm;
    bool known(int32_t idx) const { return idx >= 0 && (uint32_t)idx < n_nodes && has[idx]; }
    bool feasible(uint32_t idx) const { return idx < n_nodes && ((bits[idx >> 5] >> (idx & 31)) & 1u); }
};
static const char *kMemoKey = "yoda/b200";

// First plugin call of a cycle evaluates the pod against every node once (one device pass) and
// stores the rows in CycleState under the state lock, as collection.go:53-55 does for "Max".
// What a thread already found in a cycle: the per-node calls of one cycle come in runs on the same thread (Filter on
// thousands of nodes), and the cycle's entries never change once written, so only the first call of a run takes the
// state's read lock -- sixteen threads taking it per node would spend their time passing its cache line around.
namespace {
struct CycleSeen {
    uint64_t serial = 0;
    const void *memo = nullptr;
    bool have_max = false;
};
thread_local CycleSeen t_seen;
}  // namespace

const Yoda::Memo *Yoda::memo(framework::CycleState &state, const v1::Pod &pod, Status *st) {
    if (t_seen.serial == state.Serial() && t_seen.memo) return static_cast<const Memo *>(t_seen.memo);
    auto remember = [&](const Memo *m) {
        if (t_seen.serial != state.Serial()) t_seen = CycleSeen{state.Serial(), nullptr, false};
        t_seen.memo = m;
        return m;
    };
    state.RLock();
    const Memo *have = dynamic_cast<const Memo *>(state.Peek(kMemoKey));
    state.RUnlock();
    if (have) return remember(have);
    state.Lock();
    have = dynamic_cast<const Memo *>(state.Peek(kMemoKey));
    if (!have) {
        std::unique_lock<std::shared_mutex> lk(mu_);
        Status s = sync_locked();
        auto m = std::make_shared<Memo>();
        if (s.IsSuccess()) {
            const yoda_pod_req req = Requirement(pod);
            uint32_t n = 0;
            if (yoda_eval_batch(ctx_, &req, 1, YODA_EVAL_FULL, &m->res) != YODA_OK || yoda_node_count(ctx_, &n) != YODA_OK) {
                s = NewStatus(framework::Error, yoda_last_error());
            } else {
                n = std::min<uint32_t>(n, (uint32_t)has_scv_.size());
                m->n_nodes = n;
                m->index = index_;
                m->has.assign(has_scv_.begin(), has_scv_.begin() + n);
                if (n_rejected_)
                    for (uint32_t i = 0; i < n; ++i)
                        if (!reject_[i].empty()) m->reject.emplace(i, reject_[i]);
                m->bits.assign((n + 31) / 32 + 1, 0);
                m->raw.assign(n + 1, 0);
                m->norm.assign(n + 1, 0);
                if (n && yoda_fetch_rows(ctx_, 0, n, m->bits.data(), m->raw.data(), m->norm.data()) != YODA_OK)
                    s = NewStatus(framework::Error, yoda_last_error());
            }
        }
        if (s.IsSuccess()) {
            state.Write(kMemoKey, m);
            have = m.get();
        } else if (st) {
            *st = s;
        }
    }
    state.Unlock();
    return have ? remember(have) : nullptr;
}

// ---- plugin methods ----------------------------------------------------------------------------
// scheduler.go:76-93
Status Yoda::Filter(framework::CycleState &state, const v1::Pod &pod, const framework::NodeInfo &node) {
    const std::string &name = node.Node().Name;
    Status st;
    const Memo *m = memo(state, pod, &st);
    if (!m) return st;
    const int32_t idx = m->index_of(name);
    if (!m->known(idx)) {  // cache.Get failed (:81-84); also a node registered after this cycle's evaluation
        const std::string *why = (idx >= 0 && (uint32_t)idx < m->n_nodes) ? m->rejected((uint32_t)idx) : nullptr;
        if (why) return NewStatus(framework::Unschedulable, "Node:" + name + " Scv rejected: " + *why);
        return NewStatus(framework::Unschedulable, "Node:" + name + " Scv.core.run-linux.com \"" + name + "\" not found");
    }
    if (m->feasible((uint32_t)idx)) return NewStatus(framework::Success, "");
    return NewStatus(framework::Unschedulable, "Node:" + name);  // :92
}

// scheduler.go:95-103 + collection.go:30-57: publishes collection.Data under "Max"
std::pair<framework::PostFilterResult, Status> Yoda::PostFilter(framework::CycleState &state, const v1::Pod &pod) {
    Status st;
    const Memo *m = memo(state, pod, &st);
    if (!m) return {framework::PostFilterResult{}, st.code == framework::Success ? NewStatus(framework::Error, "evaluation failed") : st};
    auto d = std::make_shared<Data>();
    d->Value.MaxBandwidth = m->res.max_bandwidth;
    d->Value.MaxClock = m->res.max_clock;
    d->Value.MaxCore = m->res.max_core;
    d->Value.MaxFreeMemory = m->res.max_free_memory;
    d->Value.MaxPower = m->res.max_power;
    d->Value.MaxTotalMemory = m->res.max_total_memory;
    state.Lock();
    state.Write("Max", d);  // collection.go:53-55
    state.Unlock();
    return {framework::PostFilterResult{}, NewStatus(framework::Success, "")};
}

bool Yoda::Less(const framework::QueuedPodInfo &a, const framework::QueuedPodInfo &b) const {
    return GetPodPriority(a) > GetPodPriority(b);  // sort.go:8-10
}

// scheduler.go:109-130
std::pair<int64_t, Status> Yoda::Score(framework::CycleState &state, const v1::Pod &p, const std::string &nodeName) {
    if (!handle_ || !handle_->GetNodeInfo(nodeName))  // :111-114
        return {0, NewStatus(framework::Error, "getting node \"" + nodeName + "\" from Snapshot: nodeinfo not found for node name \"" + nodeName + "\"")};
    Status st;
    const Memo *m = memo(state, p, &st);
    if (!m) return {0, NewStatus(framework::Error, "Score Node Error: " + st.message)};
    const int32_t idx = m->index_of(nodeName);
    if (!m->known(idx))  // :117-122; a node registered after the cycle's evaluation is not part of it either
        return {0, NewStatus(framework::Error, "Score Node Error: Scv.core.run-linux.com \"" + nodeName + "\" not found")};
    bool have_max = t_seen.serial == state.Serial() && t_seen.have_max;  // once written it stays for the cycle
    if (!have_max) {
        state.RLock();
        have_max = dynamic_cast<const Data *>(state.Peek("Max")) != nullptr;
        state.RUnlock();
        if (t_seen.serial == state.Serial()) t_seen.have_max = have_max;
    }
    if (!have_max)  // algorithm.go:29-37: "Max" must have been written
        return {0, NewStatus(framework::Error, "Score Node Error: not found")};
    return {m->raw[idx], NewStatus(framework::Success, "")};
}

// scheduler.go:132-157.  The framework's list is the nodes feasible under ALL plugins; when it is
// Yoda's own feasible set with Yoda's scores, the device-normalized row is the answer, otherwise
// the list is rescaled on the device as given.  Panics (throws) on an empty list like the Go code.
Status Yoda::NormalizeScore(framework::CycleState &state, const v1::Pod &pod, framework::NodeScoreList &scores) {
    if (scores.empty()) throw std::out_of_range("index out of range [0] with length 0");  // scores[0], :135
    state.RLock();
    const Memo *m = dynamic_cast<const Memo *>(state.Peek(kMemoKey));
    state.RUnlock();
    bool own = m && scores.size() == m->res.feasible;
    std::vector<int32_t> idx(scores.size(), -1);
    if (own) {
        // A list in node-index order (what a framework that filters in index order hands over) is matched against the
        // feasible bits by comparing names, which is cheaper than hashing them; anything else is looked up by name.
        std::shared_lock<std::shared_mutex> lk(mu_);  // names_ (indices below m->n_nodes never change their name)
        size_t i = 0;
        bool in_order = true;
        for (uint32_t w = 0; in_order && w * 32 < m->n_nodes && i < scores.size(); ++w)
            for (uint32_t word = m->bits[w]; word && i < scores.size(); word &= word - 1) {
                const uint32_t n = w * 32 + (uint32_t)__builtin_ctz(word);
                in_order = n < m->n_nodes && names_[n] == scores[i].Name;
                if (!in_order) break;
                idx[i++] = (int32_t)n;
            }
        if (!in_order || i != scores.size()) std::fill(idx.begin(), idx.end(), -1);
    }
    for (size_t i = 0; own && i < scores.size(); ++i) {
        if (idx[i] < 0) idx[i] = m->index_of(scores[i].Name);
        // norm == 255 marks a rescale that wrapped outside [0,100]: only the exact list path returns the wrapped int64
        own = idx[i] >= 0 && m->feasible((uint32_t)idx[i]) && m->raw[idx[i]] == scores[i].Score && m->norm[idx[i]] != 255;
    }
    if (own) {
        for (size_t i = 0; i < scores.size(); ++i) scores[i].Score = m->norm[idx[i]];
        return NewStatus(framework::Success);  // the Go method returns nil
    }
    std::vector<int64_t> v(scores.size());
    for (size_t i = 0; i < scores.size(); ++i) v[i] = scores[i].Score;
    {
        std::unique_lock<std::shared_mutex> lk(mu_);
        if (yoda_normalize_scores(ctx_, v.data(), (uint32_t)v.size()) != YODA_OK) return NewStatus(framework::Error, yoda_last_error());
    }
    for (size_t i = 0; i < scores.size(); ++i) scores[i].Score = v[i];
    (void)pod;
    return NewStatus(framework::Success);
}

// ---- batch path ----------------------------------------------------------------------------------
Status Yoda::EvalBatch(const std::vector<v1::Pod> &pods, bool full, std::vector<yoda_pod_result> *out) {
    std::vector<yoda_pod_req> reqs(pods.size());
    for (size_t i = 0; i < pods.size(); ++i) reqs[i] = Requirement(pods[i]);
    out->assign(pods.size(), yoda_pod_result{});
    std::unique_lock<std::shared_mutex> lk(mu_);
    Status s = sync_locked();
    if (!s.IsSuccess()) return s;
    if (yoda_eval_batch(ctx_, reqs.data(), (uint32_t)reqs.size(), full ? YODA_EVAL_FULL : YODA_EVAL_DECIDE, out->data()) != YODA_OK)
        return NewStatus(framework::Error, yoda_last_error());
    return NewStatus(framework::Success);
}
Status Yoda::EvalSequential(const std::vector<v1::Pod> &pods, std::vector<yoda_pod_result> *out, std::vector<uint32_t> *order_out) {
    std::vector<yoda_pod_req> reqs(pods.size());
    std::vector<int64_t> pri(pods.size());
    for (size_t i = 0; i < pods.size(); ++i) {
        reqs[i] = Requirement(pods[i]);
        pri[i] = GetPodPriority(framework::QueuedPodInfo{pods[i]});
    }
    std::vector<uint32_t> order(pods.size());
    out->assign(pods.size(), yoda_pod_result{});
    if (pods.empty()) return NewStatus(framework::Success);
    std::unique_lock<std::shared_mutex> lk(mu_);
    Status s = sync_locked();
    if (!s.IsSuccess()) return s;
    if (yoda_queue_order(ctx_, pri.data(), (uint32_t)pri.size(), order.data()) != YODA_OK ||
        yoda_eval_sequential(ctx_, reqs.data(), (uint32_t)reqs.size(), order.data(), out->data()) != YODA_OK)
        return NewStatus(framework::Error, yoda_last_error());
    if (order_out) *order_out = order;
    return NewStatus(framework::Success);
}
int Yoda::FetchRow(uint32_t pod, int kind, void *dst, size_t bytes) {
    std::unique_lock<std::shared_mutex> lk(mu_);
    return yoda_fetch_row(ctx_, pod, kind, dst, bytes);
}
}  // namespace yoda
