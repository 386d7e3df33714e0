/*
 * kgwe_oracle.cpp — CPU restatement of KGWE's scheduler scoring path.
 * TEST INFRASTRUCTURE ONLY (see kgwe_oracle.h): the checker for the CUDA engine and
 * the timed CPU baseline.  Never linked into or called from the product path.
 *
 * PARITY UNPINNED BY REFERENCE TESTS (the reference has none, and no Go toolchain
 * exists here): pinned by source reading + the hand-derived KATs of SURVEY.md 8c.
 *
 * Two builds of the same algorithm from one source:
 *   namespace faithful — follows src/scheduler/scheduler.go:182-707 function by
 *     function with the reference's data structures and asymptotics: string UUIDs, a
 *     hash-map ledger that is scanned in full by getAvailableGPUs / calculateBalanceScore,
 *     adjacency maps rebuilt per (pod,node).  This is the timed "reference" CPU path.
 *   namespace indexed  — the same results on dense arrays (per-node masks), fast
 *     enough to check 10k x 10k.  Written independently of the CUDA kernels: it
 *     recomputes every aggregate from scratch in the reference's order.
 *
 * Canonical determinisation (each is an outcome Go could produce):
 * nodes visited in index order; final pick = max TotalScore, lowest index on ties;
 * NUMA groups in ascending NUMANode; PCIe groups in ascending interned switch id.
 *
 * Extensions (no counterpart in the reference, default off): E1 exhaustive subsets,
 * E3 min-bandwidth filter.  They are marked EXTENSION below.
 */
#include "kgwe_oracle.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace {

typedef long long goint;   // Go `int` on amd64

const char* kArchNames[] = {"", "Volta", "Turing", "Ampere", "Hopper", "Ada", "Blackwell"};
const char* kMigNames[KGWE_MIG_PROFILES] = {"1g.5gb", "1g.10gb", "2g.10gb", "2g.20gb", "3g.20gb",
                                            "3g.40gb", "4g.20gb", "4g.40gb", "7g.40gb", "7g.80gb"};

std::string arch_name(unsigned a) {
    if (a < 7) return kArchNames[a];
    char b[32]; snprintf(b, sizeof b, "arch-%u", a); return b;
}
std::string mig_name(unsigned p) {   // p = 1..10, or KGWE_MIG_UNKNOWN
    if (p >= 1 && p <= KGWE_MIG_PROFILES) return kMigNames[p - 1];
    return "unknown-profile";
}

uint32_t pack_order(const std::vector<int>& idx) {
    uint32_t o = 0;
    for (size_t p = 0; p < idx.size() && p < 8; ++p) o |= (uint32_t)(idx[p] & 7) << (3 * p);
    o |= (uint32_t)(idx.size() & 15) << 24;
    return o;
}

goint binom(int n, int k) {
    if (k < 0 || k > n) return 0;
    goint r = 1;
    for (int i = 1; i <= k; ++i) r = r * (n - k + i) / i;
    return r;
}

struct Weights { goint wt, wr, wb, bonus; double norm; };

Weights weights_of(const kgwe_config* c) {
    Weights w{40, 35, 25, 10, 900.0};
    if (c) { w.wt = c->topology_weight; w.wr = c->resource_weight; w.wb = c->balance_weight;
             w.bonus = c->hint_bonus; w.norm = c->nvlink_norm_gbps; }
    return w;
}

// One evaluation's result in index form (what both builds emit).
struct Eval {
    bool eligible = false, kept = false, hint = false, fallback = false, minbw = false;
    goint total = 0, topo = 0, res = 0, bal = 0;
    std::vector<int> gpus;     // SelectedGPUs, reference order
    goint candidates = 1;
};

kgwe_node_score to_record(const Eval& e) {
    kgwe_node_score r; memset(&r, 0, sizeof r);
    if (!e.eligible) return r;
    r.total = (int32_t)e.total; r.topology = (int16_t)e.topo; r.resource = (int16_t)e.res; r.balance = (int16_t)e.bal;
    uint8_t m = 0; for (int g : e.gpus) m |= (uint8_t)(1u << g);
    r.gpu_mask = m;
    r.flags = KGWE_F_ELIGIBLE | (e.kept ? KGWE_F_KEPT : 0) | (e.hint ? KGWE_F_HINT : 0) |
              (e.fallback ? KGWE_F_FALLBACK : 0) | (e.minbw ? KGWE_F_MINBW : 0);
    r.order = pack_order(e.gpus);
    return r;
}

kgwe_placement to_placement(uint32_t node_idx, const Eval& e, double est) {
    kgwe_placement p; memset(&p, 0, sizeof p);
    kgwe_node_score r = to_record(e);
    p.node_idx = node_idx; p.total = r.total; p.topology = r.topology; p.resource = r.resource;
    p.balance = r.balance; p.gpu_mask = r.gpu_mask; p.flags = r.flags; p.order = r.order;
    p.est_bandwidth_gbps = est;
    return p;
}

kgwe_placement no_placement() {
    kgwe_placement p; memset(&p, 0, sizeof p); p.node_idx = KGWE_NO_NODE; return p;
}

bool hint_bit(const uint8_t* hint, uint32_t n_nodes, uint32_t p, uint32_t n) {
    if (!hint) return false;
    size_t row = (n_nodes + 7) / 8;
    return (hint[(size_t)p * row + n / 8] >> (n & 7)) & 1;
}

// ===================================================================================
namespace faithful {

// Data model: the fields the scorer reads, src/discovery/types.go:11-47,116-164,182-209.
struct NVLinkInfo { double BandwidthGBps; bool IsActive; };
struct MIGInstance { bool InUse; std::string ProfileName; };
struct GPUDevice {
    std::string UUID;
    std::string Architecture;
    uint64_t FreeBytes;
    double GPUPercent;
    std::string HealthStatus;
    std::unordered_map<std::string, NVLinkInfo> NVLinkConnections;
    double PCIeBandwidthGBps;
    std::string SwitchID;
    int NUMANode;
    bool hasMIGConfig, MIGEnabled;
    std::vector<MIGInstance> CurrentInstances;
    int index;        // position in node.GPUs (output conversion only)
    int switchOrder;  // canonicalisation aid: iteration rank of this GPU's switch group
};
struct NodeTopology { std::string NodeName; std::vector<GPUDevice*> GPUs; uint32_t index; };
struct ClusterTopology {
    std::vector<NodeTopology*> order;                          // canonical visit order
    std::unordered_map<std::string, NodeTopology*> Nodes;      // types.go:338
};
struct GPUAllocation { std::string WorkloadUID, NodeName; std::vector<std::string> GPUUUIDs; };   // scheduler.go:68-75
struct MIGRequirements { std::string Profile; goint Count; };
struct GPUWorkload {
    std::string UID; goint GPUCount; goint MinMemoryGB; double MinBandwidthGBps;
    std::string TopologyPreference; bool hasMIG; MIGRequirements MIG; std::string GPUArchitecture;
    bool exhaustive;
};
struct PlacementHint { std::vector<std::string> PreferredNodes; };
struct NodeScore {   // types.go:295-319
    std::string NodeName; goint TotalScore = 0, TopologyScore = 0, ResourceScore = 0, BalanceScore = 0;
    std::vector<std::string> SelectedGPUs; bool hint = false, fallback = false; goint candidates = 1;
    NodeTopology* node = nullptr;
};
struct Decision { std::string SelectedNode; std::vector<std::string> SelectedGPUs; goint Score; double EstimatedBandwidth; };

bool contains(const std::vector<std::string>& s, const std::string& item) {   // scheduler.go:822-829
    for (const auto& x : s) if (x == item) return true;
    return false;
}

struct Scheduler {
    Weights cfg;
    ClusterTopology* topo;
    std::unordered_map<std::string, GPUAllocation> allocations;   // scheduler.go:29

    // scheduler.go:581-603
    std::vector<GPUDevice*> getAvailableGPUs(NodeTopology* node) {
        std::unordered_map<std::string, bool> allocatedGPUs;
        for (auto& kv : allocations) {
            if (kv.second.NodeName == node->NodeName)
                for (auto& uuid : kv.second.GPUUUIDs) allocatedGPUs[uuid] = true;
        }
        std::vector<GPUDevice*> available;
        for (GPUDevice* gpu : node->GPUs)
            if (!allocatedGPUs.count(gpu->UUID) && gpu->HealthStatus == "Healthy") available.push_back(gpu);
        return available;
    }

    // scheduler.go:606-622
    goint getMIGAvailability(NodeTopology* node, const MIGRequirements& req) {
        goint count = 0;
        for (GPUDevice* gpu : node->GPUs) {
            if (!gpu->hasMIGConfig || !gpu->MIGEnabled) continue;
            for (auto& inst : gpu->CurrentInstances)
                if (!inst.InUse && inst.ProfileName == req.Profile) count++;
        }
        return count;
    }

    // scheduler.go:206-241
    bool isNodeEligible(NodeTopology* node, const GPUWorkload& w) {
        auto availableGPUs = getAvailableGPUs(node);
        if ((goint)availableGPUs.size() < w.GPUCount) return false;
        if (!w.GPUArchitecture.empty()) {
            bool hasArch = false;
            for (GPUDevice* gpu : availableGPUs)
                if (gpu->Architecture == w.GPUArchitecture) { hasArch = true; break; }
            if (!hasArch) return false;
        }
        if (w.hasMIG) {
            if (getMIGAvailability(node, w.MIG) < w.MIG.Count) return false;
        }
        return true;
    }

    // scheduler.go:376-435
    std::pair<std::vector<std::string>, double> findBestNVLinkGroup(
        std::unordered_map<std::string, std::unordered_map<std::string, double>>& graph,
        const std::vector<GPUDevice*>& availableGPUs, goint size) {
        if (size <= 1 && !availableGPUs.empty()) return {{availableGPUs[0]->UUID}, 0.0};
        std::vector<std::string> bestGroup; double bestBandwidth = 0.0;
        for (GPUDevice* startGPU : availableGPUs) {
            std::vector<std::string> group{startGPU->UUID};
            double totalBandwidth = 0.0;
            while ((goint)group.size() < size) {
                std::string bestCandidate; double bestCandidateBW = 0.0; bool have = false;
                for (GPUDevice* candidate : availableGPUs) {
                    if (contains(group, candidate->UUID)) continue;
                    double aggBW = 0.0; bool fullyConnected = true;
                    auto& row = graph[candidate->UUID];
                    for (auto& member : group) {
                        auto it = row.find(member);
                        if (it != row.end()) aggBW += it->second; else fullyConnected = false;
                    }
                    if (fullyConnected && aggBW > bestCandidateBW) { bestCandidate = candidate->UUID; bestCandidateBW = aggBW; have = true; }
                }
                if (!have) break;
                group.push_back(bestCandidate);
                totalBandwidth += bestCandidateBW;
            }
            if ((goint)group.size() >= size && totalBandwidth > bestBandwidth) {
                bestGroup.assign(group.begin(), group.begin() + size);
                bestBandwidth = totalBandwidth;
            }
        }
        return {bestGroup, bestBandwidth};
    }

    // EXTENSION E1 (no reference counterpart): all C(A,k) subsets of the available GPUs,
    // subset valid when every ordered pair has an active link; members join in ascending
    // index, total accumulated as in the greedy (agg per joiner, then total += agg);
    // strictly-greater keeps the lowest mask on ties.
    std::pair<std::vector<std::string>, double> findBestNVLinkGroupExhaustive(
        std::unordered_map<std::string, std::unordered_map<std::string, double>>& graph,
        const std::vector<GPUDevice*>& availableGPUs, goint size) {
        std::vector<std::string> bestGroup; double bestBandwidth = 0.0;
        int A = (int)availableGPUs.size();
        for (unsigned sub = 0; sub < (1u << A); ++sub) {
            if (__builtin_popcount(sub) != size) continue;
            std::vector<GPUDevice*> g;
            for (int i = 0; i < A; ++i) if (sub >> i & 1) g.push_back(availableGPUs[i]);
            bool ok = true; double total = 0.0;
            for (size_t j = 0; j < g.size() && ok; ++j) {
                double agg = 0.0;
                for (size_t i = 0; i < j; ++i) {
                    auto a = graph[g[j]->UUID].find(g[i]->UUID);
                    auto b = graph[g[i]->UUID].find(g[j]->UUID);
                    if (a == graph[g[j]->UUID].end() || b == graph[g[i]->UUID].end()) { ok = false; break; }
                    agg += a->second;
                }
                if (j > 0) total += agg;
            }
            if (ok && total > bestBandwidth) {
                bestGroup.clear(); for (auto* d : g) bestGroup.push_back(d->UUID);
                bestBandwidth = total;
            }
        }
        return {bestGroup, bestBandwidth};
    }

    // scheduler.go:336-373
    std::pair<goint, std::vector<std::string>> scoreNVLinkTopology(
        NodeTopology*, const std::vector<GPUDevice*>& availableGPUs, goint required, bool exhaustive, bool* fallback) {
        if ((goint)availableGPUs.size() < required) return {0, {}};
        std::unordered_map<std::string, std::unordered_map<std::string, double>> nvlinkGraph;
        for (GPUDevice* gpu : availableGPUs) {
            auto& row = nvlinkGraph[gpu->UUID];
            for (auto& kv : gpu->NVLinkConnections)
                if (kv.second.IsActive) row[kv.first] = kv.second.BandwidthGBps;
        }
        auto best = exhaustive ? findBestNVLinkGroupExhaustive(nvlinkGraph, availableGPUs, required)
                               : findBestNVLinkGroup(nvlinkGraph, availableGPUs, required);
        if ((goint)best.first.size() < required) {
            std::vector<std::string> gpus;
            for (goint i = 0; i < required && i < (goint)availableGPUs.size(); ++i) gpus.push_back(availableGPUs[i]->UUID);
            *fallback = true;
            return {30, gpus};
        }
        double maxBandwidth = cfg.norm * (double)(required * (required - 1) / 2);
        double bandwidthRatio = best.second / maxBandwidth;
        goint score = (goint)(50 + 50 * bandwidthRatio);
        return {score, best.first};
    }

    // scheduler.go:438-472, groups visited in ascending NUMANode (canonical)
    std::pair<goint, std::vector<std::string>> scoreNUMATopology(
        NodeTopology*, const std::vector<GPUDevice*>& availableGPUs, goint required, bool* fallback) {
        std::map<int, std::vector<GPUDevice*>> numaGroups;
        for (GPUDevice* gpu : availableGPUs) numaGroups[gpu->NUMANode].push_back(gpu);
        for (auto& kv : numaGroups) {
            if ((goint)kv.second.size() >= required) {
                std::vector<std::string> selected;
                for (goint i = 0; i < required; ++i) selected.push_back(kv.second[i]->UUID);
                return {90, selected};
            }
        }
        std::vector<std::string> selected;
        for (auto& kv : numaGroups)
            for (GPUDevice* gpu : kv.second) {
                selected.push_back(gpu->UUID);
                if ((goint)selected.size() >= required) { *fallback = true; return {50, selected}; }
            }
        return {0, {}};
    }

    // scheduler.go:475-513, groups visited in ascending interned switch id (canonical)
    std::pair<goint, std::vector<std::string>> scorePCIeTopology(
        NodeTopology*, const std::vector<GPUDevice*>& availableGPUs, goint required, bool* fallback) {
        std::unordered_map<std::string, std::vector<GPUDevice*>> switchGroups;
        std::map<int, std::string> visit;
        for (GPUDevice* gpu : availableGPUs) {
            std::string switchID = gpu->SwitchID;
            if (switchID.empty()) switchID = "default";
            switchGroups[switchID].push_back(gpu);
            visit[gpu->switchOrder] = switchID;
        }
        for (auto& v : visit) {
            auto& gpus = switchGroups[v.second];
            if ((goint)gpus.size() >= required) {
                std::vector<std::string> selected;
                for (goint i = 0; i < required; ++i) selected.push_back(gpus[i]->UUID);
                return {80, selected};
            }
        }
        std::vector<std::string> selected;
        for (auto& v : visit)
            for (GPUDevice* gpu : switchGroups[v.second]) {
                selected.push_back(gpu->UUID);
                if ((goint)selected.size() >= required) { *fallback = true; return {40, selected}; }
            }
        return {0, {}};
    }

    // scheduler.go:303-333
    std::pair<goint, std::vector<std::string>> calculateTopologyScore(
        NodeTopology* node, const std::vector<GPUDevice*>& availableGPUs, const GPUWorkload& w, bool* fallback, goint* candidates) {
        goint required = w.GPUCount;
        *candidates = 1;
        if (required <= 1) {
            if (!availableGPUs.empty()) return {100, {availableGPUs[0]->UUID}};
            return {0, {}};
        }
        if (w.TopologyPreference == "NVLinkRequired" || w.TopologyPreference == "NVLinkOptimal") {
            *candidates = w.exhaustive ? binom((int)availableGPUs.size(), (int)required) : (goint)availableGPUs.size();
            return scoreNVLinkTopology(node, availableGPUs, required, w.exhaustive, fallback);
        }
        if (w.TopologyPreference == "SameNUMA") return scoreNUMATopology(node, availableGPUs, required, fallback);
        if (w.TopologyPreference == "SamePCIeSwitch") return scorePCIeTopology(node, availableGPUs, required, fallback);
        std::vector<std::string> gpus;
        for (goint i = 0; i < required && i < (goint)availableGPUs.size(); ++i) gpus.push_back(availableGPUs[i]->UUID);
        return {50, gpus};
    }

    // scheduler.go:516-553
    goint calculateResourceScore(NodeTopology*, const std::vector<GPUDevice*>& availableGPUs, const GPUWorkload& w) {
        if (availableGPUs.empty()) return 0;
        goint score = 50;
        uint64_t totalFreeMemory = 0;
        for (GPUDevice* gpu : availableGPUs) totalFreeMemory += gpu->FreeBytes;
        uint64_t requiredMemory = (uint64_t)w.MinMemoryGB * 1024 * 1024 * 1024 * (uint64_t)w.GPUCount;
        if (totalFreeMemory >= requiredMemory * 2) score += 25;
        else if (totalFreeMemory >= requiredMemory) score += 10;
        double avgUtilization = 0.0;
        for (GPUDevice* gpu : availableGPUs) avgUtilization += gpu->GPUPercent;
        avgUtilization /= (double)availableGPUs.size();
        if (avgUtilization < 30) score += 25;
        else if (avgUtilization < 60) score += 10;
        return std::min<goint>(100, score);
    }

    // scheduler.go:556-578
    goint calculateBalanceScore(NodeTopology* node) {
        goint nodeAllocations = 0;
        for (auto& kv : allocations) if (kv.second.NodeName == node->NodeName) nodeAllocations++;
        goint totalGPUs = (goint)node->GPUs.size();
        if (totalGPUs == 0) return 0;
        double usageRatio = (double)nodeAllocations / (double)totalGPUs;
        goint score = (goint)(100 * (1 - usageRatio));
        return std::max<goint>(0, score);
    }

    // scheduler.go:244-300
    NodeScore scoreNode(NodeTopology* node, const GPUWorkload& w, const PlacementHint* hint) {
        NodeScore score; score.NodeName = node->NodeName; score.node = node;
        auto availableGPUs = getAvailableGPUs(node);
        auto ts = calculateTopologyScore(node, availableGPUs, w, &score.fallback, &score.candidates);
        score.TopologyScore = ts.first; score.SelectedGPUs = ts.second;
        score.ResourceScore = calculateResourceScore(node, availableGPUs, w);
        score.BalanceScore = calculateBalanceScore(node);
        goint hintBonus = 0;
        if (hint) for (auto& pn : hint->PreferredNodes) if (pn == node->NodeName) { hintBonus = cfg.bonus; score.hint = true; break; }
        score.TotalScore = (score.TopologyScore * cfg.wt + score.ResourceScore * cfg.wr + score.BalanceScore * cfg.wb) / 100;
        score.TotalScore += hintBonus;
        return score;
    }

    // scheduler.go:656-692 (GetNodeTopology: discovery.go:199-204)
    double estimateBandwidth(const NodeScore& score) {
        if (score.SelectedGPUs.size() <= 1) return 0;
        auto it = topo->Nodes.find(score.NodeName);
        if (it == topo->Nodes.end()) return 0;
        NodeTopology* nodeTopology = it->second;
        double totalBandwidth = 0.0; goint pairs = 0;
        for (size_t i = 0; i < score.SelectedGPUs.size(); ++i)
            for (size_t j = 0; j < score.SelectedGPUs.size(); ++j) {
                if (i >= j) continue;
                for (GPUDevice* gpu : nodeTopology->GPUs) {
                    if (gpu->UUID == score.SelectedGPUs[i]) {
                        auto nv = gpu->NVLinkConnections.find(score.SelectedGPUs[j]);
                        if (nv != gpu->NVLinkConnections.end()) totalBandwidth += nv->second.BandwidthGBps;
                        else totalBandwidth += gpu->PCIeBandwidthGBps;
                        pairs++;
                    }
                }
            }
        if (pairs == 0) return 0;
        return totalBandwidth / (double)pairs;
    }

    // One (pod,node) pass: isNodeEligible + scoreNode + the `TotalScore > 0` keep rule
    // of scoreNodes (scheduler.go:189-200), plus EXTENSION E3.
    bool evaluate(NodeTopology* node, const GPUWorkload& w, const PlacementHint* hint, NodeScore* out, bool* eligible, bool* minbw) {
        *minbw = false;
        *eligible = isNodeEligible(node, w);
        if (!*eligible) return false;
        *out = scoreNode(node, w, hint);
        bool kept = out->TotalScore > 0;
        // EXTENSION E3: drop a node whose chosen GPUs cannot give the requested inter-GPU bandwidth.
        if (kept && w.MinBandwidthGBps > 0 && w.GPUCount >= 2 && estimateBandwidth(*out) < w.MinBandwidthGBps) { *minbw = true; kept = false; }
        return kept;
    }

    // scheduler.go:625-653
    bool tryScheduleOnNode(const NodeScore& ns, Decision* d) {
        for (auto& uuid : ns.SelectedGPUs)
            for (auto& kv : allocations)
                if (contains(kv.second.GPUUUIDs, uuid)) return false;
        d->SelectedNode = ns.NodeName; d->SelectedGPUs = ns.SelectedGPUs; d->Score = ns.TotalScore;
        d->EstimatedBandwidth = estimateBandwidth(ns);
        return true;
    }

    // scheduler.go:695-707
    void recordAllocation(const GPUWorkload& w, const Decision& d) {
        allocations[w.UID] = GPUAllocation{w.UID, d.SelectedNode, d.SelectedGPUs};
    }
};

std::string uuid_of(uint32_t node, int gpu) { char b[48]; snprintf(b, sizeof b, "GPU-%08x-%d", node, gpu); return b; }
std::string node_name(uint32_t node) { char b[32]; snprintf(b, sizeof b, "node-%06u", node); return b; }
std::string wl_name(uint64_t id) { char b[40]; snprintf(b, sizeof b, "wl-%llu", (unsigned long long)id); return b; }

struct World {
    std::vector<GPUDevice> gpus; std::vector<NodeTopology> nodes; ClusterTopology topo;
    std::unordered_map<std::string, int> gpuIndex;    // uuid -> index in node (output conversion only)

    World(const kgwe_node_image* imgs, uint32_t n) {
        gpus.reserve((size_t)n * KGWE_MAX_GPUS); nodes.resize(n);
        for (uint32_t i = 0; i < n; ++i) {
            const kgwe_node_image& im = imgs[i];
            NodeTopology& nt = nodes[i]; nt.NodeName = node_name(i); nt.index = i;
            int ng = std::min<int>(im.n_gpus, KGWE_MAX_GPUS);
            for (int g = 0; g < ng; ++g) {
                GPUDevice d;
                d.UUID = uuid_of(i, g); d.Architecture = arch_name(im.arch[g]); d.FreeBytes = im.free_bytes[g];
                d.GPUPercent = im.util[g]; d.HealthStatus = (im.healthy_mask >> g & 1) ? "Healthy" : "Unhealthy";
                for (int j = 0; j < ng; ++j)
                    if (j != g && (im.nvl_present[g] >> j & 1))
                        d.NVLinkConnections[uuid_of(i, j)] = NVLinkInfo{im.nvl_bw[g][j], (bool)(im.nvl_active[g] >> j & 1)};
                d.PCIeBandwidthGBps = im.pcie_bw[g];
                char sw[16]; snprintf(sw, sizeof sw, "sw%03u", im.pcie_switch[g]); d.SwitchID = sw;
                d.switchOrder = im.pcie_switch[g];
                d.NUMANode = im.numa[g];
                d.hasMIGConfig = false; d.MIGEnabled = false; d.index = g;
                if (g == 0) {   // the image keeps per-node free-instance counts; hang them on GPU 0
                    bool any = false;
                    for (int p = 0; p < KGWE_MIG_PROFILES; ++p)
                        for (int c = 0; c < im.mig_free[p]; ++c) { d.CurrentInstances.push_back(MIGInstance{false, kMigNames[p]}); any = true; }
                    if (any) { d.hasMIGConfig = true; d.MIGEnabled = true; d.CurrentInstances.push_back(MIGInstance{true, kMigNames[1]}); }
                }
                gpus.push_back(std::move(d));
            }
        }
        size_t at = 0;
        for (uint32_t i = 0; i < n; ++i) {
            int ng = std::min<int>(imgs[i].n_gpus, KGWE_MAX_GPUS);
            for (int g = 0; g < ng; ++g) { nodes[i].GPUs.push_back(&gpus[at]); gpuIndex[gpus[at].UUID] = g; ++at; }
            topo.order.push_back(&nodes[i]); topo.Nodes[nodes[i].NodeName] = &nodes[i];
        }
    }
};

GPUWorkload workload_of(const kgwe_pod_req& p) {
    GPUWorkload w;
    w.UID = wl_name(p.wl_id); w.GPUCount = p.gpu_count; w.MinMemoryGB = p.min_memory_gb; w.MinBandwidthGBps = p.min_bandwidth_gbps;
    switch (p.topology_pref) {
        case KGWE_TOPO_NVLINK_OPTIMAL: w.TopologyPreference = "NVLinkOptimal"; break;
        case KGWE_TOPO_NVLINK_REQUIRED: w.TopologyPreference = "NVLinkRequired"; break;
        case KGWE_TOPO_SAME_NUMA: w.TopologyPreference = "SameNUMA"; break;
        case KGWE_TOPO_SAME_PCIE_SWITCH: w.TopologyPreference = "SamePCIeSwitch"; break;
        default: w.TopologyPreference = "None";
    }
    w.hasMIG = p.mig_profile != KGWE_MIG_NONE; w.MIG.Profile = mig_name(p.mig_profile); w.MIG.Count = p.mig_count;
    w.GPUArchitecture = p.arch == KGWE_ARCH_ANY ? "" : (p.arch == 255 ? "no-such-arch" : arch_name(p.arch));
    w.exhaustive = (p.flags & KGWE_POD_EXHAUSTIVE) != 0;
    return w;
}

Eval eval_of(World& W, bool eligible, bool kept, bool minbw, const NodeScore& ns) {
    Eval e; e.eligible = eligible;
    if (!eligible) return e;
    e.kept = kept; e.minbw = minbw; e.hint = ns.hint; e.fallback = ns.fallback;
    e.total = ns.TotalScore; e.topo = ns.TopologyScore; e.res = ns.ResourceScore; e.bal = ns.BalanceScore;
    for (auto& u : ns.SelectedGPUs) e.gpus.push_back(W.gpuIndex[u]);
    e.candidates = ns.candidates;
    return e;
}

void seed_ledger(Scheduler& S, const kgwe_oracle_alloc* ledger, uint32_t n) {
    for (uint32_t i = 0; i < n; ++i) {
        GPUAllocation a; a.WorkloadUID = wl_name(ledger[i].wl_id); a.NodeName = node_name(ledger[i].node_idx);
        for (int g = 0; g < KGWE_MAX_GPUS; ++g) if (ledger[i].gpu_mask >> g & 1) a.GPUUUIDs.push_back(uuid_of(ledger[i].node_idx, g));
        S.allocations[a.WorkloadUID] = a;
    }
}

}  // namespace faithful

// ===================================================================================
namespace indexed {

struct NodeState { uint8_t allocated; uint32_t n_allocs; };

inline int avail_list(const kgwe_node_image& im, const NodeState& st, int* out) {   // scheduler.go:581-603
    int ng = std::min<int>(im.n_gpus, KGWE_MAX_GPUS), a = 0;
    for (int g = 0; g < ng; ++g)
        if (!(st.allocated >> g & 1) && (im.healthy_mask >> g & 1)) out[a++] = g;
    return a;
}

inline bool link_active(const kgwe_node_image& im, int from, int to) {
    return (im.nvl_present[from] >> to & 1) && (im.nvl_active[from] >> to & 1);
}

double est_bandwidth(const kgwe_node_image& im, const std::vector<int>& sel) {   // scheduler.go:656-692
    if (sel.size() <= 1) return 0;
    double total = 0.0; goint pairs = 0;
    for (size_t i = 0; i < sel.size(); ++i)
        for (size_t j = i + 1; j < sel.size(); ++j) {
            if (im.nvl_present[sel[i]] >> sel[j] & 1) total += im.nvl_bw[sel[i]][sel[j]];
            else total += im.pcie_bw[sel[i]];
            pairs++;
        }
    if (pairs == 0) return 0;
    return total / (double)pairs;
}

Eval evaluate(const Weights& cfg, const kgwe_node_image& im, const NodeState& st, const kgwe_pod_req& p, bool hinted) {
    Eval e;
    int av[KGWE_MAX_GPUS]; int A = avail_list(im, st, av);
    goint k = p.gpu_count;
    // isNodeEligible, scheduler.go:206-241
    if (A < k) return e;
    if (p.arch != KGWE_ARCH_ANY) {
        bool has = false;
        for (int i = 0; i < A; ++i) if (im.arch[av[i]] == p.arch) { has = true; break; }
        if (!has) return e;
    }
    if (p.mig_profile != KGWE_MIG_NONE) {
        goint have = (p.mig_profile >= 1 && p.mig_profile <= KGWE_MIG_PROFILES) ? im.mig_free[p.mig_profile - 1] : 0;
        if (have < p.mig_count) return e;
    }
    e.eligible = true;
    // calculateTopologyScore, scheduler.go:303-333
    if (k <= 1) {
        if (A > 0) { e.topo = 100; e.gpus = {av[0]}; } else e.topo = 0;
    } else if (p.topology_pref == KGWE_TOPO_NVLINK_OPTIMAL || p.topology_pref == KGWE_TOPO_NVLINK_REQUIRED) {
        bool exhaustive = (p.flags & KGWE_POD_EXHAUSTIVE) != 0;
        std::vector<int> best; double bestBW = 0.0;
        if (!exhaustive) {   // scheduler.go:376-435
            e.candidates = A;
            for (int s = 0; s < A; ++s) {
                std::vector<int> group{av[s]}; double total = 0.0;
                while ((goint)group.size() < k) {
                    int bc = -1; double bcBW = 0.0;
                    for (int c = 0; c < A; ++c) {
                        int cand = av[c];
                        if (std::find(group.begin(), group.end(), cand) != group.end()) continue;
                        double agg = 0.0; bool full = true;
                        for (int m : group) { if (link_active(im, cand, m)) agg += im.nvl_bw[cand][m]; else full = false; }
                        if (full && agg > bcBW) { bc = cand; bcBW = agg; }
                    }
                    if (bc < 0) break;
                    group.push_back(bc); total += bcBW;
                }
                if ((goint)group.size() >= k && total > bestBW) { best.assign(group.begin(), group.begin() + k); bestBW = total; }
            }
        } else {             // EXTENSION E1
            e.candidates = binom(A, (int)k);
            for (unsigned sub = 0; sub < (1u << A); ++sub) {
                if (__builtin_popcount(sub) != k) continue;
                int g[KGWE_MAX_GPUS], n = 0;
                for (int i = 0; i < A; ++i) if (sub >> i & 1) g[n++] = av[i];
                bool ok = true; double total = 0.0;
                for (int j = 0; j < n && ok; ++j) {
                    double agg = 0.0;
                    for (int i = 0; i < j; ++i) {
                        if (!link_active(im, g[j], g[i]) || !link_active(im, g[i], g[j])) { ok = false; break; }
                        agg += im.nvl_bw[g[j]][g[i]];
                    }
                    if (j > 0) total += agg;
                }
                if (ok && total > bestBW) { best.assign(g, g + n); bestBW = total; }
            }
        }
        if ((goint)best.size() < k) {   // scheduler.go:358-365
            e.topo = 30; e.fallback = true;
            for (goint i = 0; i < k && i < A; ++i) e.gpus.push_back(av[i]);
        } else {                        // scheduler.go:368-370
            double maxBandwidth = cfg.norm * (double)(k * (k - 1) / 2);
            double ratio = bestBW / maxBandwidth;
            e.topo = (goint)(50 + 50 * ratio);
            e.gpus = best;
        }
    } else if (p.topology_pref == KGWE_TOPO_SAME_NUMA || p.topology_pref == KGWE_TOPO_SAME_PCIE_SWITCH) {
        // scheduler.go:438-472 / 475-513: group by key, groups in ascending key
        bool numa = p.topology_pref == KGWE_TOPO_SAME_NUMA;
        std::map<int, std::vector<int>> groups;
        for (int i = 0; i < A; ++i) groups[numa ? (int)im.numa[av[i]] : (int)im.pcie_switch[av[i]]].push_back(av[i]);
        bool done = false;
        for (auto& kv : groups)
            if ((goint)kv.second.size() >= k) { e.topo = numa ? 90 : 80; e.gpus.assign(kv.second.begin(), kv.second.begin() + k); done = true; break; }
        if (!done) {
            for (auto& kv : groups) { for (int g : kv.second) { e.gpus.push_back(g); if ((goint)e.gpus.size() >= k) { done = true; break; } } if (done) break; }
            if (done) { e.topo = numa ? 50 : 40; e.fallback = true; } else { e.topo = 0; e.gpus.clear(); }
        }
    } else {
        e.topo = 50;
        for (goint i = 0; i < k && i < A; ++i) e.gpus.push_back(av[i]);
    }
    // calculateResourceScore, scheduler.go:516-553
    if (A == 0) e.res = 0;
    else {
        goint score = 50; uint64_t totalFree = 0;
        for (int i = 0; i < A; ++i) totalFree += im.free_bytes[av[i]];
        uint64_t required = (uint64_t)p.min_memory_gb * 1024 * 1024 * 1024 * (uint64_t)k;
        if (totalFree >= required * 2) score += 25; else if (totalFree >= required) score += 10;
        double avg = 0.0;
        for (int i = 0; i < A; ++i) avg += im.util[av[i]];
        avg /= (double)A;
        if (avg < 30) score += 25; else if (avg < 60) score += 10;
        e.res = std::min<goint>(100, score);
    }
    // calculateBalanceScore, scheduler.go:556-578
    {
        goint totalGPUs = std::min<int>(im.n_gpus, KGWE_MAX_GPUS);
        if (totalGPUs == 0) e.bal = 0;
        else {
            double usage = (double)st.n_allocs / (double)totalGPUs;
            e.bal = std::max<goint>(0, (goint)(100 * (1 - usage)));
        }
    }
    // scoreNode total, scheduler.go:282-286
    e.total = (e.topo * cfg.wt + e.res * cfg.wr + e.bal * cfg.wb) / 100;
    if (hinted) { e.total += cfg.bonus; e.hint = true; }
    e.kept = e.total > 0;
    // EXTENSION E3
    if (e.kept && p.min_bandwidth_gbps > 0 && k >= 2 && est_bandwidth(im, e.gpus) < p.min_bandwidth_gbps) { e.minbw = true; e.kept = false; }
    return e;
}

std::vector<NodeState> states_from_ledger(uint32_t n_nodes, const kgwe_oracle_alloc* ledger, uint32_t n) {
    std::vector<NodeState> st(n_nodes, NodeState{0, 0});
    for (uint32_t i = 0; i < n; ++i) if (ledger[i].node_idx < n_nodes) { st[ledger[i].node_idx].allocated |= ledger[i].gpu_mask; st[ledger[i].node_idx].n_allocs++; }
    return st;
}

}  // namespace indexed

struct Ranked { goint total; uint32_t node; Eval e; double est; };

void emit_topk(std::vector<Ranked>& kept, uint32_t topk, kgwe_placement* out) {
    std::stable_sort(kept.begin(), kept.end(), [](const Ranked& a, const Ranked& b) { return a.total > b.total; });
    for (uint32_t j = 0; j < topk; ++j) out[j] = j < kept.size() ? to_placement(kept[j].node, kept[j].e, kept[j].est) : no_placement();
}

bool args_ok(const kgwe_pod_req* pods, uint32_t n_pods) {
    for (uint32_t i = 0; i < n_pods; ++i)
        if (pods[i].gpu_count < 0 || pods[i].min_memory_gb < 0 || pods[i].min_memory_gb > (1 << 20)) return false;
    return true;
}

}  // namespace

extern "C" int kgwe_oracle_score_batch(const kgwe_config* cfg, const kgwe_node_image* nodes, uint32_t n_nodes,
                                       const kgwe_oracle_alloc* ledger, uint32_t n_ledger,
                                       const kgwe_pod_req* pods, uint32_t n_pods, const uint8_t* hint,
                                       kgwe_node_score* out_scores, uint32_t topk, kgwe_placement* out_topk,
                                       int mode, int threads, kgwe_oracle_counters* counters) {
    if (!nodes || !pods || topk > KGWE_MAX_TOPK || !args_ok(pods, n_pods)) return KGWE_ERR_INVALID_ARG;
    Weights W = weights_of(cfg);
    uint32_t base = cfg ? cfg->node_base : 0;
    if (threads < 1) threads = 1;
    std::vector<uint64_t> evs(threads, 0), cands(threads, 0);
    faithful::World* world = mode == KGWE_ORACLE_FAITHFUL ? new faithful::World(nodes, n_nodes) : nullptr;
    std::vector<indexed::NodeState> states = indexed::states_from_ledger(n_nodes, ledger, n_ledger);
    std::vector<faithful::Scheduler> scheds;
    if (world) {
        scheds.resize(threads);
        for (auto& s : scheds) { s.cfg = W; s.topo = &world->topo; faithful::seed_ledger(s, ledger, n_ledger); }
    }
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&](int tid) {
        for (uint32_t p = tid; p < n_pods; p += threads) {
            std::vector<Ranked> kept;
            faithful::GPUWorkload w; faithful::PlacementHint ph;
            if (world) {
                w = faithful::workload_of(pods[p]);
                if (hint) for (uint32_t n = 0; n < n_nodes; ++n) if (hint_bit(hint, n_nodes, p, n)) ph.PreferredNodes.push_back(faithful::node_name(n));
            }
            for (uint32_t n = 0; n < n_nodes; ++n) {
                Eval e; double est = 0;
                if (world) {
                    faithful::NodeScore ns; bool el, mb;
                    bool k = scheds[tid].evaluate(world->topo.order[n], w, hint ? &ph : nullptr, &ns, &el, &mb);
                    e = faithful::eval_of(*world, el, k, mb, ns);
                    if (k && topk) est = scheds[tid].estimateBandwidth(ns);
                } else {
                    e = indexed::evaluate(W, nodes[n], states[n], pods[p], hint_bit(hint, n_nodes, p, n));
                    if (e.kept && topk) est = indexed::est_bandwidth(nodes[n], e.gpus);
                }
                evs[tid]++; cands[tid] += (uint64_t)e.candidates;
                if (out_scores) out_scores[(size_t)p * n_nodes + n] = to_record(e);
                if (e.kept && topk) kept.push_back(Ranked{e.total, base + n, e, est});
            }
            if (topk && out_topk) emit_topk(kept, topk, out_topk + (size_t)p * topk);
        }
    };
    if (threads == 1) work(0);
    else { std::vector<std::thread> th; for (int t = 0; t < threads; ++t) th.emplace_back(work, t); for (auto& t : th) t.join(); }
    auto t1 = std::chrono::steady_clock::now();
    if (counters) {
        counters->evaluations = 0; counters->candidates = 0;
        for (int t = 0; t < threads; ++t) { counters->evaluations += evs[t]; counters->candidates += cands[t]; }
        counters->seconds = std::chrono::duration<double>(t1 - t0).count();
    }
    delete world;
    return KGWE_OK;
}

extern "C" int kgwe_oracle_schedule_queue(const kgwe_config* cfg, const kgwe_node_image* nodes, uint32_t n_nodes,
                                          const kgwe_oracle_alloc* ledger, uint32_t n_ledger,
                                          const kgwe_pod_req* pods, uint32_t n_pods, const uint8_t* hint,
                                          kgwe_placement* out, kgwe_oracle_alloc* ledger_out, uint32_t* n_ledger_out,
                                          int mode, kgwe_oracle_counters* counters) {
    if (!nodes || !pods || !out || !args_ok(pods, n_pods)) return KGWE_ERR_INVALID_ARG;
    Weights W = weights_of(cfg);
    uint32_t base = cfg ? cfg->node_base : 0;
    uint64_t evs = 0, cands = 0;
    // index-form ledger kept by both modes for ledger_out (and as the indexed state)
    std::vector<kgwe_oracle_alloc> led(ledger, ledger + n_ledger);
    auto find_led = [&](uint64_t id) { for (size_t i = 0; i < led.size(); ++i) if (led[i].wl_id == id) return (long)i; return -1L; };
    faithful::World* world = mode == KGWE_ORACLE_FAITHFUL ? new faithful::World(nodes, n_nodes) : nullptr;
    faithful::Scheduler S; S.cfg = W;
    if (world) { S.topo = &world->topo; faithful::seed_ledger(S, ledger, n_ledger); }
    std::vector<indexed::NodeState> states = indexed::states_from_ledger(n_nodes, ledger, n_ledger);
    auto t0 = std::chrono::steady_clock::now();
    for (uint32_t p = 0; p < n_pods; ++p) {
        // Schedule, scheduler.go:114-179 (optimizer == nil unless a hint bitmap is given)
        bool found = false; goint bestTotal = 0; uint32_t bestNode = 0; Eval bestEval; double est = 0;
        if (world) {
            faithful::GPUWorkload w = faithful::workload_of(pods[p]);
            faithful::PlacementHint ph;
            if (hint) for (uint32_t n = 0; n < n_nodes; ++n) if (hint_bit(hint, n_nodes, p, n)) ph.PreferredNodes.push_back(faithful::node_name(n));
            std::vector<faithful::NodeScore> nodeScores;   // scoreNodes, scheduler.go:182-203
            for (uint32_t n = 0; n < n_nodes; ++n) {
                faithful::NodeScore ns; bool el, mb;
                bool k = S.evaluate(world->topo.order[n], w, hint ? &ph : nullptr, &ns, &el, &mb);
                evs++; cands += el ? (uint64_t)ns.candidates : 1;
                if (k) nodeScores.push_back(ns);
            }
            // sort desc (scheduler.go:144-146), canonically stable => lowest index first among ties
            std::stable_sort(nodeScores.begin(), nodeScores.end(), [](const faithful::NodeScore& a, const faithful::NodeScore& b) { return a.TotalScore > b.TotalScore; });
            for (auto& ns : nodeScores) {
                faithful::Decision d;
                if (S.tryScheduleOnNode(ns, &d)) {
                    S.recordAllocation(w, d);
                    found = true; bestTotal = ns.TotalScore; bestNode = ns.node->index; est = d.EstimatedBandwidth;
                    bestEval = faithful::eval_of(*world, true, true, false, ns);
                    break;
                }
            }
        } else {
            for (uint32_t n = 0; n < n_nodes; ++n) {
                Eval e = indexed::evaluate(W, nodes[n], states[n], pods[p], hint_bit(hint, n_nodes, p, n));
                evs++; cands += (uint64_t)e.candidates;
                if (e.kept && (!found || e.total > bestTotal)) { found = true; bestTotal = e.total; bestNode = n; bestEval = e; }
            }
            if (found) est = indexed::est_bandwidth(nodes[bestNode], bestEval.gpus);
        }
        if (!found) { out[p] = no_placement(); continue; }
        out[p] = to_placement(base + bestNode, bestEval, est);
        // recordAllocation: same UID overwrites (scheduler.go:699)
        uint8_t mask = out[p].gpu_mask;
        long old = find_led(pods[p].wl_id);
        if (old >= 0) {
            led.erase(led.begin() + old);
        }
        kgwe_oracle_alloc a; memset(&a, 0, sizeof a); a.wl_id = pods[p].wl_id; a.node_idx = bestNode; a.gpu_mask = mask;
        led.push_back(a);
        if (!world) {
            if (old >= 0) states = indexed::states_from_ledger(n_nodes, led.data(), (uint32_t)led.size());
            else { states[bestNode].allocated |= mask; states[bestNode].n_allocs++; }
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    if (ledger_out && n_ledger_out) { memcpy(ledger_out, led.data(), led.size() * sizeof(kgwe_oracle_alloc)); *n_ledger_out = (uint32_t)led.size(); }
    if (counters) { counters->evaluations = evs; counters->candidates = cands; counters->seconds = std::chrono::duration<double>(t1 - t0).count(); }
    delete world;
    return KGWE_OK;
}

// EXTENSION E5 (no reference counterpart beyond the GangSchedulingGroup type, types.go:415-444)
extern "C" int kgwe_oracle_schedule_gangs(const kgwe_config* cfg, const kgwe_node_image* nodes, uint32_t n_nodes,
                                          const kgwe_oracle_alloc* ledger, uint32_t n_ledger,
                                          const kgwe_pod_req* pods, uint32_t n_pods, const uint32_t* gang_of, const uint8_t* hint,
                                          kgwe_placement* out, kgwe_oracle_alloc* ledger_out, uint32_t* n_ledger_out, int mode) {
    if (!nodes || !pods || !out || !gang_of) return KGWE_ERR_INVALID_ARG;
    std::vector<kgwe_oracle_alloc> led(ledger, ledger + n_ledger);
    const size_t row = (n_nodes + 7) / 8;
    uint32_t p0 = 0;
    while (p0 < n_pods) {
        uint32_t p1 = p0 + 1;
        while (p1 < n_pods && gang_of[p1] == gang_of[p0]) ++p1;
        const uint32_t np = p1 - p0;
        std::vector<kgwe_oracle_alloc> next(led.size() + np);
        uint32_t n_next = 0;
        int rc = kgwe_oracle_schedule_queue(cfg, nodes, n_nodes, led.data(), (uint32_t)led.size(), pods + p0, np,
                                            hint ? hint + (size_t)p0 * row : nullptr, out + p0, next.data(), &n_next, mode, nullptr);
        if (rc != KGWE_OK) return rc;
        bool all = true;
        for (uint32_t i = 0; i < np; ++i) if (out[p0 + i].node_idx == KGWE_NO_NODE) all = false;
        if (all) { next.resize(n_next); led.swap(next); }
        else for (uint32_t i = 0; i < np; ++i) {            // roll the whole gang back
            const bool had = out[p0 + i].node_idx != KGWE_NO_NODE;
            out[p0 + i] = no_placement();
            out[p0 + i].flags = had ? KGWE_F_GANG_ROLLBACK : 0;
        }
        p0 = p1;
    }
    if (ledger_out && n_ledger_out) { memcpy(ledger_out, led.data(), led.size() * sizeof(kgwe_oracle_alloc)); *n_ledger_out = (uint32_t)led.size(); }
    return KGWE_OK;
}

// EXTENSION E4
extern "C" int kgwe_oracle_placement_terms(const kgwe_config* cfg, const kgwe_node_image* nodes, uint32_t n_nodes,
                                           const kgwe_placement* pls, uint32_t n, kgwe_side_terms* out) {
    if (!nodes || !pls || !out) return KGWE_ERR_INVALID_ARG;
    const uint32_t base = cfg ? cfg->node_base : 0;
    for (uint32_t i = 0; i < n; ++i) {
        out[i].bisection_gbps = 0.0; out[i].ring_bottleneck_gbps = 0.0;
        if (pls[i].node_idx == KGWE_NO_NODE || pls[i].node_idx - base >= n_nodes) continue;
        const kgwe_node_image& im = nodes[pls[i].node_idx - base];
        std::vector<int> g;
        for (int b = 0; b < KGWE_MAX_GPUS; ++b) if (pls[i].gpu_mask >> b & 1) g.push_back(b);
        const int k = (int)g.size();
        if (k < 2) continue;
        auto w = [&](int a, int b) -> double {
            if (!indexed::link_active(im, g[a], g[b]) || !indexed::link_active(im, g[b], g[a])) return 0.0;
            return std::min(im.nvl_bw[g[a]][g[b]], im.nvl_bw[g[b]][g[a]]);
        };
        double best = 1e300;
        for (unsigned t = 0; t < (1u << k); ++t) {
            if (__builtin_popcount(t) != k / 2) continue;
            double cut = 0.0;
            for (int a = 0; a < k; ++a) if (t >> a & 1) for (int b = 0; b < k; ++b) if (!(t >> b & 1)) cut += w(a, b);
            best = std::min(best, cut);
        }
        std::vector<int> perm;
        for (int q = 1; q < k; ++q) perm.push_back(q);
        double ring = -1.0;
        do {
            double mn = w(0, perm[0]);
            for (int q = 0; q + 1 < k - 1; ++q) mn = std::min(mn, w(perm[q], perm[q + 1]));
            mn = std::min(mn, w(perm[k - 2], 0));
            ring = std::max(ring, mn);
        } while (std::next_permutation(perm.begin(), perm.end()));
        out[i].bisection_gbps = best; out[i].ring_bottleneck_gbps = ring;
    }
    return KGWE_OK;
}

// EXTENSION E2 — MIG placement tables (NVIDIA MIG placement rules; instance maxima agree with
// the reference's profile table, src/discovery/types.go:233-239: 7 / 3 / 2 / 1 / 1).
namespace mig {
struct Placement { int geo, start, size; unsigned mask; };
const int kSize[5] = {1, 2, 4, 4, 8};
std::vector<Placement> legal() {
    std::vector<Placement> v;
    for (int s = 0; s <= 6; ++s) v.push_back({0, s, 1, 1u << s});
    for (int s = 0; s <= 4; s += 2) v.push_back({1, s, 2, 3u << s});
    for (int s = 0; s <= 4; s += 4) v.push_back({2, s, 4, 15u << s});
    v.push_back({3, 0, 4, 15u});
    v.push_back({4, 0, 8, 255u});
    return v;
}
}  // namespace mig

extern "C" int kgwe_oracle_mig_pack(kgwe_mig_gpu* gpus, uint32_t n_gpus, const uint8_t* profiles, uint32_t n_reqs, kgwe_mig_assignment* out) {
    if ((n_gpus && !gpus) || (n_reqs && (!profiles || !out))) return KGWE_ERR_INVALID_ARG;
    const std::vector<mig::Placement> all = mig::legal();
    for (uint32_t r = 0; r < n_reqs; ++r) {
        const int p = profiles[r];
        kgwe_mig_assignment a; memset(&a, 0, sizeof a); a.gpu_idx = KGWE_NO_NODE; a.profile = (uint8_t)p;
        if (p >= 1 && p <= KGWE_MIG_PROFILES) {
            const int geo = (p - 1) / 2, cls = (p - 1) % 2;
            int bestCost = 1 << 30;
            for (uint32_t g = 0; g < n_gpus; ++g) {
                if (!gpus[g].enabled || (gpus[g].cls & 1) != cls) continue;
                for (const auto& pl : all) {
                    if (pl.geo != geo || (gpus[g].occ & pl.mask)) continue;
                    int cost = 0;
                    for (const auto& q : all) if (!(gpus[g].occ & q.mask) && (q.mask & pl.mask)) ++cost;
                    if (cost < bestCost) { bestCost = cost; a.gpu_idx = g; a.start = (uint8_t)pl.start; a.size = (uint8_t)pl.size; a.cost = (uint8_t)cost; }
                }
            }
            if (a.gpu_idx != KGWE_NO_NODE) gpus[a.gpu_idx].occ |= (uint8_t)(((1u << a.size) - 1u) << a.start);
        }
        out[r] = a;
    }
    return KGWE_OK;
}

namespace mig {
// How many GPUs each geometry of class c gets out of `count`: shares of the positive targets of that class, rounded down,
// the GPUs left over handed out one each in order of decreasing fractional part (ties: lower geometry first).  Written
// from the spec in include/kgwe_engine.h with a sort, on purpose differently from the engine's host code.
std::vector<uint32_t> split(uint32_t count, const double* target, int c) {
    std::vector<uint32_t> n(5, 0);
    double total = 0.0;
    for (int q = 0; q < 5; ++q) if (target[2 * q + c] > 0.0) total += target[2 * q + c];
    if (!(total > 0.0) || count == 0) return n;
    struct Rem { double frac; int q; };
    std::vector<Rem> rems;
    uint32_t given = 0;
    for (int q = 0; q < 5; ++q) {
        if (!(target[2 * q + c] > 0.0)) continue;
        const double share = (target[2 * q + c] / total) * (double)count;
        n[q] = (uint32_t)share;
        given += n[q];
        rems.push_back({share - (double)n[q], q});
    }
    std::stable_sort(rems.begin(), rems.end(), [](const Rem& a, const Rem& b) { return a.frac > b.frac; });      // stable: ties keep geometry order
    for (size_t i = 0; i < rems.size() && given < count; ++i, ++given) n[rems[i].q]++;
    return n;
}
const int kStarts[5] = {7, 3, 2, 1, 1}, kStride[5] = {1, 2, 4, 0, 0};
}  // namespace mig

extern "C" int kgwe_oracle_mig_rebalance(const kgwe_mig_gpu* gpus, uint32_t n_gpus, const double* target,
                                         kgwe_mig_op* ops, uint32_t ops_cap, uint32_t* n_ops, uint32_t* supply) {
    if ((n_gpus && !gpus) || !target || !n_ops || !supply) return KGWE_ERR_INVALID_ARG;
    std::vector<kgwe_mig_op> all;
    for (int p = 0; p < KGWE_MIG_PROFILES; ++p) supply[p] = 0;
    for (int c = 0; c < 2; ++c) {
        std::vector<uint32_t> idle;
        for (uint32_t g = 0; g < n_gpus; ++g) if (gpus[g].enabled && gpus[g].occ == 0 && (gpus[g].cls & 1) == c) idle.push_back(g);
        const std::vector<uint32_t> n = mig::split((uint32_t)idle.size(), target, c);
        size_t at = 0;
        for (int q = 0; q < 5; ++q)
            for (uint32_t i = 0; i < n[q]; ++i, ++at)
                for (int sidx = 0; sidx < mig::kStarts[q]; ++sidx) {
                    kgwe_mig_op op; memset(&op, 0, sizeof op);
                    op.gpu_idx = idle[at]; op.profile = (uint8_t)(2 * q + 1 + c); op.start = (uint8_t)(sidx * mig::kStride[q]); op.size = (uint8_t)mig::kSize[q];
                    all.push_back(op); supply[2 * q + c]++;
                }
    }
    *n_ops = (uint32_t)all.size();
    for (uint32_t i = 0; i < all.size() && i < ops_cap; ++i) ops[i] = all[i];
    return KGWE_OK;
}

// The reconfiguration pass with destroy ops (spec: kgwe_mig_reconfigure in include/kgwe_engine.h), by plain loops over the
// 14 legal placements in bit order.
extern "C" int kgwe_oracle_mig_reconfigure(const kgwe_mig_gpu_state* gpus, uint32_t n_gpus, const double* target,
                                           kgwe_mig_op* ops, uint32_t ops_cap, uint32_t* n_ops, uint32_t* supply) {
    if ((n_gpus && !gpus) || !target || !n_ops || !supply) return KGWE_ERR_INVALID_ARG;
    const std::vector<mig::Placement> legal = mig::legal();          // index = placement bit
    std::vector<kgwe_mig_op> all;
    for (int p = 0; p < KGWE_MIG_PROFILES; ++p) supply[p] = 0;
    for (uint32_t g = 0; g < n_gpus; ++g) {                          // GPUs with something running keep what they have
        if (!gpus[g].enabled || gpus[g].in_use == 0) continue;
        for (int b = 0; b < 14; ++b)
            if ((gpus[g].created >> b & 1) && !(gpus[g].in_use >> b & 1)) supply[2 * legal[b].geo + (gpus[g].cls & 1)]++;
    }
    for (int c = 0; c < 2; ++c) {
        std::vector<uint32_t> free_gpus;
        for (uint32_t g = 0; g < n_gpus; ++g) if (gpus[g].enabled && gpus[g].in_use == 0 && (gpus[g].cls & 1) == c) free_gpus.push_back(g);
        const std::vector<uint32_t> n = mig::split((uint32_t)free_gpus.size(), target, c);
        size_t at = 0;
        for (int q = 0; q < 5; ++q)
            for (uint32_t i = 0; i < n[q]; ++i, ++at) {
                const uint32_t g = free_gpus[at];
                for (int b = 0; b < 14; ++b)                         // whatever exists and is not wanted goes
                    if ((gpus[g].created >> b & 1) && legal[b].geo != q) {
                        kgwe_mig_op op; memset(&op, 0, sizeof op);
                        op.gpu_idx = g; op.profile = (uint8_t)(2 * legal[b].geo + 1 + c); op.start = (uint8_t)legal[b].start; op.size = (uint8_t)legal[b].size; op.pad = KGWE_MIG_OP_DESTROY;
                        all.push_back(op);
                    }
                for (int b = 0; b < 14; ++b)                         // whatever is wanted and does not exist yet comes
                    if (legal[b].geo == q && !(gpus[g].created >> b & 1)) {
                        kgwe_mig_op op; memset(&op, 0, sizeof op);
                        op.gpu_idx = g; op.profile = (uint8_t)(2 * q + 1 + c); op.start = (uint8_t)legal[b].start; op.size = (uint8_t)legal[b].size; op.pad = KGWE_MIG_OP_CREATE;
                        all.push_back(op);
                    }
                supply[2 * q + c] += (uint32_t)mig::kStarts[q];
            }
    }
    *n_ops = (uint32_t)all.size();
    for (uint32_t i = 0; i < all.size() && i < ops_cap; ++i) ops[i] = all[i];
    return KGWE_OK;
}

// ---- DiscoveryService.GetTopologyHint, src/discovery/discovery.go:222-247 ----------------------
namespace hint {
struct Scored { goint score; std::vector<int> gpus; };

// getAvailableGPUs, discovery.go:436-460
std::vector<int> getAvailableGPUs(const kgwe_node_image& im) {
    std::vector<int> available;
    const int ng = std::min<int>(im.n_gpus, KGWE_MAX_GPUS);
    for (int g = 0; g < ng; ++g) {
        if (!(im.healthy_mask >> g & 1)) continue;
        if (im.mig_enabled_mask >> g & 1) { if (im.mig_has_free_mask >> g & 1) available.push_back(g); }
        else if (im.util[g] < 90) available.push_back(g);
    }
    return available;
}

// findNVLinkGroups, discovery.go:462-488 (peers of a GPU visited in ascending GPU index: canonical)
std::vector<std::vector<int>> findNVLinkGroups(const kgwe_node_image& im, const std::vector<int>& gpus, goint size) {
    std::vector<std::vector<int>> groups;
    const int ng = std::min<int>(im.n_gpus, KGWE_MAX_GPUS);
    for (int gpu : gpus) {
        std::vector<int> group{gpu};
        for (int peer = 0; peer < ng; ++peer) {
            if (!(im.nvl_present[gpu] >> peer & 1) || peer == gpu) continue;
            if ((goint)group.size() >= size) break;
            if (std::find(gpus.begin(), gpus.end(), peer) != gpus.end()) group.push_back(peer);
        }
        if ((goint)group.size() >= size) { group.resize(size); groups.push_back(group); }
    }
    return groups;
}

// scoreNodeForRequirements, discovery.go:378-434
Scored scoreNodeForRequirements(const kgwe_node_image& im, const kgwe_hint_req& req) {
    std::vector<int> availableGPUs = getAvailableGPUs(im);
    if ((goint)availableGPUs.size() < req.count) return {0, {}};
    std::vector<int> filteredGPUs;
    for (int g : availableGPUs) {
        goint memGB = (goint)(im.free_bytes[g] / (1024ull * 1024 * 1024));
        if (memGB >= req.min_memory_gb) filteredGPUs.push_back(g);
    }
    if ((goint)filteredGPUs.size() < req.count) return {0, {}};
    goint score = 50;
    std::vector<int> selected;
    if (req.count > 1 && req.prefer_nvlink) {
        auto groups = findNVLinkGroups(im, filteredGPUs, req.count);
        if (!groups.empty()) { score += 30; selected = groups[0]; }
    }
    if (selected.empty())
        for (goint i = 0; i < req.count && i < (goint)filteredGPUs.size(); ++i) selected.push_back(filteredGPUs[i]);
    if (req.prefer_same_numa) {           // areOnSameNUMA, discovery.go:490-504
        bool same = true; int numaNode = -1000;
        for (int g : selected) { if (numaNode == -1000) numaNode = im.numa[g]; else if (numaNode != im.numa[g]) same = false; }
        if (same) score += 10;
    }
    if (req.arch != KGWE_ARCH_ANY)
        for (int g : selected) if (im.arch[g] == req.arch) score += 5;
    return {score, selected};
}
}  // namespace hint

extern "C" int kgwe_oracle_topology_hints(const kgwe_config* cfg, const kgwe_node_image* nodes, uint32_t n_nodes,
                                          const kgwe_hint_req* reqs, uint32_t n_reqs, kgwe_topology_hint* out) {
    if (!nodes || (n_reqs && (!reqs || !out))) return KGWE_ERR_INVALID_ARG;
    const uint32_t base = cfg ? cfg->node_base : 0;
    for (uint32_t r = 0; r < n_reqs; ++r) {
        kgwe_topology_hint t; memset(&t, 0, sizeof t); t.node_idx = KGWE_NO_NODE;
        goint bestScore = 0; std::vector<int> bestGpus; uint32_t bestNode = 0;
        for (uint32_t n = 0; n < n_nodes; ++n) {                       // discovery.go:231-238, nodes in index order
            hint::Scored s = hint::scoreNodeForRequirements(nodes[n], reqs[r]);
            if (s.score > bestScore) { bestScore = s.score; bestGpus = s.gpus; bestNode = n; }
        }
        if (bestScore > 0) {
            t.node_idx = base + bestNode; t.score = (int32_t)bestScore; t.order = pack_order(bestGpus);
            for (int g : bestGpus) t.gpu_mask |= (uint8_t)(1u << g);
            t.bandwidth_estimate = indexed::est_bandwidth(nodes[bestNode], bestGpus);     // discovery.go:506-541 (same formula as scheduler.go:656-692)
        }
        out[r] = t;
    }
    return KGWE_OK;
}

// PrometheusExporter.calculateTopologyScore, src/monitoring/prometheus_exporter.go:517-539
extern "C" int kgwe_oracle_topology_gauges(const kgwe_node_image* nodes, uint32_t n_nodes, int32_t* out) {
    if (n_nodes && (!nodes || !out)) return KGWE_ERR_INVALID_ARG;
    for (uint32_t n = 0; n < n_nodes; ++n) {
        if (nodes[n].n_gpus == 0) { out[n] = 0; continue; }
        goint score = 50;
        if (nodes[n].node_flags & KGWE_NODE_NVSWITCH) score += 30;
        if (nodes[n].node_flags & KGWE_NODE_ANY_NVLINK) score += 20;
        out[n] = (int32_t)std::min<goint>(100, score);
    }
    return KGWE_OK;
}
