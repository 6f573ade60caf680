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
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>

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

static uint64_t fnv(uint64_t h, uint64_t v) {
    for (int i = 0; i < 8; ++i) { h ^= (v >> (8 * i)) & 0xFF; h *= 0x100000001B3ull; }
    return h;
}

int main(int argc, char **argv) {
    uint64_t N = 2000, P = 256, pod_seed = synth::SEED + 2, cycles = 16, seq = 0;
    bool checksum = false;
    const char *dump = nullptr;
    for (int i = 1; i < argc; ++i) {
        auto next = [&]() { return (i + 1 < argc) ? strtoull(argv[++i], nullptr, 10) : 0ull; };
        if (!strcmp(argv[i], "--nodes")) N = next();
        else if (!strcmp(argv[i], "--pods")) P = next();
        else if (!strcmp(argv[i], "--pod-seed")) pod_seed = next();
        else if (!strcmp(argv[i], "--cycles")) cycles = next();
        else if (!strcmp(argv[i], "--seq")) seq = next();
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
    opt.max_nodes = (uint32_t)std::max<uint64_t>(N, 1);
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

    // ---- (2) framework path ---------------------------------------------------------------------------
    int mismatches = 0;
    cycles = std::min<uint64_t>(cycles, P);
    t0 = Clock::now();
    for (uint64_t p = 0; p < cycles; ++p) {
        framework::CycleState state;
        const v1::Pod &pod = pods[p];
        std::vector<std::string> feasible;
        for (const auto &nd : nodes)  // findNodesThatPassFilters
            if (plugin->Filter(state, pod, *handle.GetNodeInfo(nd.scv.Name)).IsSuccess()) feasible.push_back(nd.scv.Name);
        auto pf = plugin->PostFilter(state, pod);  // max-collection hook
        if (!pf.second.IsSuccess()) { mismatches++; continue; }
        int32_t chosen = -1;
        if (!feasible.empty()) {
            framework::NodeScoreList scores;
            for (const auto &nm : feasible) {  // RunScorePlugins
                auto sc = plugin->Score(state, pod, nm);
                if (!sc.second.IsSuccess()) { mismatches++; break; }
                scores.push_back({nm, sc.first});
            }
            plugin->NormalizeScore(state, pod, scores);
            int64_t best = -1;
            for (const auto &s : scores)  // selectHost with the deterministic tie-break of the harness
                if (s.Score > best) { best = s.Score; chosen = plugin->NodeIndex(s.Name); }
        }
        auto mx = std::dynamic_pointer_cast<yoda::Data>(state.Read("Max"));
        if (chosen != batch[p].chosen || feasible.size() != batch[p].feasible || !mx || mx->Value.MaxFreeMemory != batch[p].max_free_memory)
            mismatches++;
    }
    const double cycle_ms = cycles ? ms_since(t0) / (double)cycles : 0.0;

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
           "\"framework_cycle_ms\": %.3f, \"cycles_checked\": %llu, \"seq_pods\": %llu, \"seq_ms\": %.3f, \"seq_mismatches\": %d, \"mismatches\": %d, \"path\": %u}\n",
           (unsigned long long)N, (unsigned long long)P, load_ms, batch_ms, tm.total_ms, (double)N * (double)P / (batch_ms * 1e-3), cycle_ms,
           (unsigned long long)cycles, (unsigned long long)seq, seq_ms, seq_mismatches, mismatches, tm.path);
    return mismatches ? 1 : 0;
}
