// harness.cpp -- the benchmark harness the north star asks cmd/scheduler to gain: it drives the
// plugin through a fake framework handle (no cluster is available offline).
//
//   yoda_harness --checksum [--nodes N --pods P --pod-seed S]     CPU only: digests of the synthetic data
//   yoda_harness [--nodes N --pods P --pod-seed S --cycles K --dump FILE]
//
// GPU run: builds the synthetic cluster, feeds it to the plugin as informer events, then
//   (1) batch path: every pending pod in one device pass (Yoda::EvalBatch), timed;
//   (2) framework path for the first K pods: the call sequence kube-scheduler makes per pod --
//       Filter on every node, PostFilter (the reference's max-collection hook), Score on every
//       feasible node, NormalizeScore, selectHost (here: lowest index among the top scores);
// and checks (2) against (1).  Exit code 0 = identical.  One JSON line on stdout.
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <atomic>
#include <fstream>
#include <functional>
#include <thread>

#include "synth.hpp"
#include "yoda_plugin.hpp"

using Clock = std::chrono::steady_clock;
static double ms_since(Clock::time_point t0) { return std::chrono::duration<double, std::milli>(Clock::now() - t0).count(); }

struct FakeHandle : framework::Handle {
    std::unordered_map<std::string, framework::NodeInfo> infos;
    const framework::NodeInfo *GetNodeInfo(const std::string &name) const override {
        auto it = infos.find(name);
        return it == infos.end() ? nullptr : &it->second;
    }
};

// A fixed set of worker threads handing out pieces of 64 indices, like parallelize.Until with its chunking.
class Workers {
   public:
    explicit Workers(unsigned n) {
        for (unsigned t = 0; t < n; ++t) pool_.emplace_back([this]() { loop(); });
    }
    ~Workers() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
            ++gen_;
        }
        cv_.notify_all();
        for (auto &th : pool_) th.join();
    }
    void run(size_t n, const std::function<void(size_t)> &f) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            n_ = n;
            f_ = &f;
            next_.store(0);
            left_ = (unsigned)pool_.size();
            ++gen_;
        }
        cv_.notify_all();
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [&]() { return left_ == 0; });
    }

   private:
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&]() { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
            }
            for (size_t i = next_.fetch_add(64); i < n_; i = next_.fetch_add(64))
                for (size_t k = i; k < std::min(n_, i + 64); ++k) (*f_)(k);
            std::lock_guard<std::mutex> lk(mu_);
            if (--left_ == 0) done_.notify_one();
        }
    }
    std::vector<std::thread> pool_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    std::atomic<size_t> next_{0};
    size_t n_ = 0;
    const std::function<void(size_t)> *f_ = nullptr;
    unsigned left_ = 0;
    uint64_t gen_ = 0;
    bool stop_ = false;
};

static uint64_t fnv(uint64_t h, uint64_t v) {
    for (int i = 0; i < 8; ++i) { h ^= (v >> (8 * i)) & 0xFF; h *= 0x100000001B3ull; }
    return h;
}

int main(int argc, char **argv) {
    uint64_t N = 2000, P = 256, pod_seed = synth::SEED + 2, cycles = 16, seq = 0, threads = 16;
    bool checksum = false;
    const char *dump = nullptr;
    for (int i = 1; i < argc; ++i) {
        auto next = [&]() { return (i + 1 < argc) ? strtoull(argv[++i], nullptr, 10) : 0ull; };
        if (!strcmp(argv[i], "--nodes")) N = next();
        else if (!strcmp(argv[i], "--pods")) P = next();
        else if (!strcmp(argv[i], "--pod-seed")) pod_seed = next();
        else if (!strcmp(argv[i], "--cycles")) cycles = next();
        else if (!strcmp(argv[i], "--seq")) seq = next();
        else if (!strcmp(argv[i], "--threads")) threads = std::max<uint64_t>(1, next());  // framework parallelism (16 in kube-scheduler)
        else if (!strcmp(argv[i], "--checksum")) checksum = true;
        else if (!strcmp(argv[i], "--dump") && i + 1 < argc) dump = argv[++i];
    }
    std::vector<synth::SynthNode> nodes;
    for (uint64_t n = 0; n < N; ++n) nodes.push_back(synth::make_node(n));
    std::vector<v1::Pod> pods;
    for (uint64_t p = 0; p < P; ++p) pods.push_back(synth::make_pod(p, pod_seed));

    if (checksum) {  // digest in a fixed field order, mirrored by tests/test_host_cpp.py
        uint64_t h = 0xCBF29CE484222325ull;
        for (const auto &nd : nodes) {
            h = fnv(h, nd.has_scv); h = fnv(h, nd.scv.Status.CardNumber); h = fnv(h, nd.scv.Status.CardList.size());
            h = fnv(h, nd.scv.Status.FreeMemorySum); h = fnv(h, nd.scv.Status.TotalMemorySum);
            for (const auto &c : nd.scv.Status.CardList) {
                h = fnv(h, c.Health == "Healthy"); h = fnv(h, c.FreeMemory); h = fnv(h, c.TotalMemory); h = fnv(h, c.Clock);
                h = fnv(h, c.Core); h = fnv(h, c.Bandwidth); h = fnv(h, c.Power);
            }
            for (uint64_t m : nd.pod_mem) h = fnv(h, m);
        }
        uint64_t hp = 0xCBF29CE484222325ull;
        for (const auto &pod : pods) {
            const yoda_pod_req r = yoda::Yoda::Requirement(pod);
            hp = fnv(hp, r.flags); hp = fnv(hp, r.number); hp = fnv(hp, r.memory); hp = fnv(hp, r.clock);
        }
        printf("{\"nodes\": %llu, \"pods\": %llu, \"snapshot_digest\": \"%016llx\", \"pods_digest\": \"%016llx\"}\n", (unsigned long long)N,
               (unsigned long long)P, (unsigned long long)h, (unsigned long long)hp);
        return 0;
    }

    // ---- fake cluster -> plugin ------------------------------------------------------------------
    FakeHandle handle;
    std::string err;
    yoda::Yoda::Options opt;
    opt.max_nodes = (uint32_t)std::max<uint64_t>(N, 1) + 1;  // one spare index for the late node of (2b)
    auto plugin = yoda::Yoda::New(opt, &handle, &err);
    if (!plugin) { fprintf(stderr, "yoda.New failed: %s\n", err.c_str()); return 2; }
    auto t0 = Clock::now();
    for (const auto &nd : nodes) {
        framework::NodeInfo info;
        info.node.Name = nd.scv.Name;
        for (size_t k = 0; k < nd.pod_mem.size(); ++k) {
            v1::Pod running;
            running.Name = nd.scv.Name + "-run-" + std::to_string(k);
            running.Labels["scv/memory"] = std::to_string(nd.pod_mem[k]);
            info.Pods.push_back({running});
            plugin->OnPodAdd(nd.scv.Name, running);
        }
        handle.infos[nd.scv.Name] = info;
        if (nd.has_scv) {
            auto st = plugin->OnScvUpsert(nd.scv);
            if (!st.IsSuccess()) { fprintf(stderr, "upsert: %s\n", st.message.c_str()); return 2; }
        } else {
            plugin->OnScvDelete(nd.scv.Name);  // the node exists, its Scv object does not
        }
    }
    if (!plugin->Sync().IsSuccess()) { fprintf(stderr, "sync failed: %s\n", yoda_last_error()); return 2; }
    const double load_ms = ms_since(t0);

    // ---- (1) batch path ----------------------------------------------------------------------------
    std::vector<yoda_pod_result> batch;
    auto st = plugin->EvalBatch(pods, true, &batch);  // warm-up (allocations)
    if (!st.IsSuccess()) { fprintf(stderr, "EvalBatch: %s\n", st.message.c_str()); return 2; }
    t0 = Clock::now();
    st = plugin->EvalBatch(pods, true, &batch);
    const double batch_ms = ms_since(t0);
    yoda_timing tm{};
    yoda_last_timing(plugin->ctx(), &tm);

    // ---- (2) framework path: Filter and Score from `threads` threads at once, as parallelize.Until does ----------
    int mismatches = 0;
    cycles = std::min<uint64_t>(cycles, P);
    std::vector<double> cycle_times;
    Workers workers((unsigned)threads);  // kube-scheduler keeps its 16 workers too (parallelize.Until)
    auto parallel_nodes = [&](size_t n, const std::function<void(size_t)> &f) { workers.run(n, f); };
    t0 = Clock::now();
    for (uint64_t p = 0; p < cycles; ++p) {
        const auto c0 = Clock::now();
        framework::CycleState state;
        const v1::Pod &pod = pods[p];
        std::vector<uint8_t> fit(nodes.size(), 0);
        parallel_nodes(nodes.size(), [&](size_t n) {  // findNodesThatPassFilters
            fit[n] = plugin->Filter(state, pod, *handle.GetNodeInfo(nodes[n].scv.Name)).IsSuccess();
        });
        std::vector<const std::string *> feasible;
        for (size_t n = 0; n < nodes.size(); ++n)
            if (fit[n]) feasible.push_back(&nodes[n].scv.Name);
        auto pf = plugin->PostFilter(state, pod);  // max-collection hook
        if (!pf.second.IsSuccess()) { mismatches++; continue; }
        int32_t chosen = -1;
        if (!feasible.empty()) {
            framework::NodeScoreList scores(feasible.size());
            std::atomic<int> bad{0};
            parallel_nodes(feasible.size(), [&](size_t k) {  // RunScorePlugins
                auto sc = plugin->Score(state, pod, *feasible[k]);
                if (!sc.second.IsSuccess()) bad++;
                scores[k] = {*feasible[k], sc.first};
            });
            mismatches += bad.load();
            // odd cycles hand the list over back to front: NormalizeScore must not depend on the order it comes in
            // (in node order it matches names against the feasible bits, otherwise it looks every name up)
            if (p & 1) std::reverse(scores.begin(), scores.end());
            plugin->NormalizeScore(state, pod, scores);
            if (p & 1) std::reverse(scores.begin(), scores.end());
            int64_t best = -1;
            for (const auto &s : scores)  // selectHost with the deterministic tie-break of the harness
                if (s.Score > best) { best = s.Score; chosen = plugin->NodeIndex(s.Name); }
        }
        cycle_times.push_back(ms_since(c0));
        state.Lock();
        auto mx = std::dynamic_pointer_cast<yoda::Data>(state.Read("Max"));
        state.Unlock();
        if (chosen != batch[p].chosen || feasible.size() != batch[p].feasible || !mx || mx->Value.MaxFreeMemory != batch[p].max_free_memory)
            mismatches++;
    }
    const double cycle_ms = cycles ? ms_since(t0) / (double)cycles : 0.0;
    std::sort(cycle_times.begin(), cycle_times.end());
    const double cycle_p50 = cycle_times.empty() ? 0.0 : cycle_times[cycle_times.size() / 2];
    const double cycle_p99 = cycle_times.empty() ? 0.0 : cycle_times[std::min(cycle_times.size() - 1, (size_t)(cycle_times.size() * 0.99))];

    // ---- (2b) a node whose Scv arrives between Filter and Score of a cycle is not part of that cycle: Score must
    //          answer "not found" (as the reference does for a node without Scv), never index past the memo; and
    //          a refused Scv (TotalMemorySum == 0) takes the node's previous record out of service
    if (P > 0) {
        framework::CycleState state;
        const v1::Pod &pod = pods[0];
        plugin->Filter(state, pod, *handle.GetNodeInfo(nodes[0].scv.Name));
        plugin->PostFilter(state, pod);
        scv::Scv late = nodes[0].scv;
        late.Name = "late-node";
        framework::NodeInfo li;
        li.node.Name = late.Name;
        handle.infos[late.Name] = li;
        if (N + 1 <= opt.max_nodes) {
            plugin->OnScvUpsert(late);
            auto sc = plugin->Score(state, pod, late.Name);
            if (sc.second.IsSuccess() || sc.second.message.find("not found") == std::string::npos) mismatches++;
            if (plugin->Filter(state, pod, li).IsSuccess()) mismatches++;
            plugin->OnScvDelete(late.Name);
        }
        size_t victim = 0;
        while (victim < nodes.size() && !nodes[victim].has_scv) ++victim;
        if (victim < nodes.size()) {
            scv::Scv broken = nodes[victim].scv;
            broken.Status.TotalMemorySum = 0;
            if (plugin->OnScvUpsert(broken).IsSuccess()) mismatches++;
            framework::CycleState st2;
            auto f = plugin->Filter(st2, pod, *handle.GetNodeInfo(broken.Name));
            if (f.IsSuccess() || f.message.find("rejected") == std::string::npos) mismatches++;
            plugin->OnScvUpsert(nodes[victim].scv);  // repaired
        }
        plugin->Sync();
    }

    // ---- (3) sequential commit: one device pass for the first `seq` pods against the real thing -- one
    //         framework cycle per pod with the pod assumed onto its node (OnPodAdd) before the next cycle
    int seq_mismatches = 0;
    double seq_ms = 0.0;
    seq = std::min<uint64_t>(seq, P);
    if (seq) {
        std::vector<v1::Pod> head(pods.begin(), pods.begin() + seq);
        std::vector<yoda_pod_result> sres;
        t0 = Clock::now();
        auto sst = plugin->EvalSequential(head, &sres);
        seq_ms = ms_since(t0);
        if (!sst.IsSuccess()) { fprintf(stderr, "EvalSequential: %s\n", sst.message.c_str()); return 2; }
        std::vector<std::pair<std::string, uint64_t>> assumed;
        for (uint64_t p = 0; p < seq; ++p) {
            framework::CycleState state;
            std::vector<std::string> feasible;
            for (const auto &nd : nodes)
                if (plugin->Filter(state, head[p], *handle.GetNodeInfo(nd.scv.Name)).IsSuccess()) feasible.push_back(nd.scv.Name);
            plugin->PostFilter(state, head[p]);
            int32_t chosen = -1;
            if (!feasible.empty()) {
                framework::NodeScoreList scores;
                for (const auto &nm : feasible) scores.push_back({nm, plugin->Score(state, head[p], nm).first});
                plugin->NormalizeScore(state, head[p], scores);
                int64_t best = -1;
                for (const auto &sc : scores)
                    if (sc.Score > best) { best = sc.Score; chosen = plugin->NodeIndex(sc.Name); }
            }
            if (chosen != sres[p].chosen) seq_mismatches++;
            if (chosen >= 0) {  // assume: the pod now counts towards the node's allocateMemorySum
                plugin->OnPodAdd(plugin->NodeName((uint32_t)chosen), head[p]);
                assumed.push_back({plugin->NodeName((uint32_t)chosen), p});
            }
        }
        for (const auto &a : assumed) plugin->OnPodDelete(a.first, head[a.second]);  // undo
        plugin->Sync();
        mismatches += seq_mismatches;
    }

    if (dump) {
        std::ofstream f(dump);
        for (const auto &r : batch) f << r.chosen << " " << r.feasible << " " << r.lowest << " " << r.highest << " " << r.max_free_memory << "\n";
    }
    printf("{\"nodes\": %llu, \"pods\": %llu, \"load_ms\": %.3f, \"batch_ms\": %.3f, \"batch_device_ms\": %.3f, \"evals_per_s\": %.4g, "
           "\"framework_cycle_ms\": %.3f, \"framework_cycle_p50_ms\": %.3f, \"framework_cycle_p99_ms\": %.3f, \"threads\": %llu, \"cycles_checked\": %llu, \"seq_pods\": %llu, \"seq_ms\": %.3f, \"seq_mismatches\": %d, \"mismatches\": %d, \"path\": %u}\n",
           (unsigned long long)N, (unsigned long long)P, load_ms, batch_ms, tm.total_ms, (double)N * (double)P / (batch_ms * 1e-3), cycle_ms, cycle_p50, cycle_p99,
           (unsigned long long)threads, (unsigned long long)cycles, (unsigned long long)seq, seq_ms, seq_mismatches, mismatches, tm.path);
    return mismatches ? 1 : 0;
}
