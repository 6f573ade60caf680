// synth.hpp -- C++ twin of yoda_scheduler_b200/synth.py: the synthetic SCV snapshot and pending-pod
// generator (counter-based SplitMix64; every value is a pure function of (seed, stream, i, j)).
// tests/test_host_cpp.py checks that both twins produce identical data.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "framework.hpp"

namespace synth {
constexpr uint64_t SEED = 0x59304441ull;
enum : uint64_t { S_SKU = 1, S_FREE, S_HEALTH, S_CARDNUM, S_NOSCV, S_NPODS, S_PODMEM, S_NCARDS,
                  S_P_NUM = 16, S_P_NUMV, S_P_MEM, S_P_MEMV, S_P_CLK, S_P_CLKV };
struct Sku { uint64_t total, clock, core, bandwidth, power; };
static const Sku SKU_TABLE[8] = {
    {11178, 5705, 3584, 352, 250},    {11019, 7000, 4352, 616, 260},   {16160, 877, 5120, 900, 300},
    {32510, 877, 5120, 900, 300},     {40536, 1215, 6912, 1555, 400},  {81920, 1593, 6912, 2039, 400},
    {81559, 2619, 16896, 3350, 700},  {183359, 3996, 18944, 8000, 1000}};
constexpr uint64_t BOGUS_CLOCK = 1234;

inline uint64_t sm64(uint64_t x) {
    uint64_t z = x + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline uint64_t rnd(uint64_t seed, uint64_t stream, uint64_t i, uint64_t j = 0) { return sm64(sm64(sm64(seed + stream) + i) + j); }

struct SynthNode {
    scv::Scv scv;
    bool has_scv;
    std::vector<uint64_t> pod_mem;  // scv/memory of the pods running on the node
};

inline SynthNode make_node(uint64_t idx, uint64_t seed = SEED) {
    SynthNode n;
    n.scv.Name = "node-" + std::to_string(idx);
    const Sku &sku = SKU_TABLE[rnd(seed, S_SKU, idx) % 8];
    const uint64_t odd = rnd(seed, S_NCARDS, idx);
    const uint64_t ncards = (odd % 512 == 0) ? (odd >> 16) % 9 : 8;
    uint64_t fsum = 0, tsum = 0;
    for (uint64_t g = 0; g < ncards; ++g) {
        scv::Card c;
        c.TotalMemory = sku.total;
        c.FreeMemory = (sku.total * (rnd(seed, S_FREE, idx, g) >> 32)) >> 32;
        c.Health = (rnd(seed, S_HEALTH, idx, g) % 64 != 0) ? "Healthy" : "Unhealthy";
        c.Clock = sku.clock; c.Core = sku.core; c.Bandwidth = sku.bandwidth; c.Power = sku.power;
        fsum += c.FreeMemory;
        tsum += c.TotalMemory;
        n.scv.Status.CardList.push_back(c);
    }
    const uint64_t cn = rnd(seed, S_CARDNUM, idx);
    n.scv.Status.CardNumber = (cn % 128 == 0) ? (cn >> 8) % 17 : 8;
    n.scv.Status.FreeMemorySum = fsum;
    n.scv.Status.TotalMemorySum = tsum ? tsum : 1;
    n.has_scv = rnd(seed, S_NOSCV, idx) % 256 != 0;
    const uint64_t npods = rnd(seed, S_NPODS, idx) % 9;
    for (uint64_t k = 0; k < npods; ++k) n.pod_mem.push_back((rnd(seed, S_PODMEM, idx, k) % 128 + 1) * 256);
    return n;
}

inline v1::Pod make_pod(uint64_t p, uint64_t seed) {
    v1::Pod pod;
    pod.Name = "pod-" + std::to_string(p);
    pod.Namespace = "default";
    if (rnd(seed, S_P_NUM, p) % 100 >= 50) {
        const uint64_t v = rnd(seed, S_P_NUMV, p);
        static const uint64_t tab[4] = {1, 2, 4, 8};
        const uint64_t num = ((v >> 20) % 100 == 0) ? ((v & 1) ? 9 : 0) : tab[(v >> 8) % 4];
        pod.Labels["scv/number"] = std::to_string(num);
    }
    if (rnd(seed, S_P_MEM, p) % 100 >= 20) pod.Labels["scv/memory"] = std::to_string(rnd(seed, S_P_MEMV, p) % 90000 + 1);
    if (rnd(seed, S_P_CLK, p) % 100 >= 85) {
        const uint64_t v = rnd(seed, S_P_CLKV, p);
        pod.Labels["scv/clock"] = std::to_string(((v >> 20) % 100 == 0) ? BOGUS_CLOCK : SKU_TABLE[(v >> 8) % 8].clock);
    }
    return pod;
}
}  // namespace synth
