// A small cluster in the reference's own struct form (discovery types), built by hand the way a
// Go test would build it: string UUIDs, NVLinkConnections maps, switch ids, NUMA nodes, MIG instances.
#pragma once
#include <cstdio>
#include <string>
#include <vector>

#include "../../k8s-gpu-workload-enhancer_b200/host/kgwe_host.hpp"

namespace fixture {
using namespace kgwe_host;

inline std::string uuid(int node, int gpu) { char b[64]; snprintf(b, sizeof b, "GPU-%04d-%d-aaaa-bbbb", node, gpu); return b; }

inline uint64_t lcg(uint64_t& s) { s = s * 6364136223846793005ULL + 1442695040888963407ULL; return s >> 33; }

// kind 0: H100 DGX full mesh 900; 1: A100 full mesh 600, 32 GB/s PCIe; 2: B200 full mesh 1800; 3: A100 cube mesh
inline NodeTopology make_node(int idx, int kind, uint64_t& rng) {
    NodeTopology n;
    char name[32]; snprintf(name, sizeof name, "node-%03d", idx);
    n.NodeName = name;
    n.NVSwitchPresent = kind != 3;
    const double bw = kind == 0 ? 900.0 : kind == 2 ? 1800.0 : 600.0;
    const char* arch = kind == 0 ? "Hopper" : kind == 2 ? "Blackwell" : "Ampere";
    const int total_gib = kind == 2 ? 180 : 80;
    const int numa_ids[2] = {7, -1};                       // deliberately not 0/1 and not ascending: -1 sorts first
    const char* switches[4] = {"", "pcie-sw-B", "pcie-sw-A", "zz"};
    for (int g = 0; g < 8; ++g) {
        GPUDevice d;
        d.UUID = uuid(idx, g); d.Index = g; d.Architecture = arch;
        d.MemoryFreeBytes = (uint64_t)(total_gib / 2 + (int)(lcg(rng) % (total_gib / 2 + 1))) << 30;
        d.UtilizationGPUPercent = (double)(lcg(rng) % 200) / 2.0;
        d.HealthStatus = (lcg(rng) % 50 == 0) ? "Degraded" : "Healthy";
        for (int p = 0; p < 8; ++p) {
            if (p == g) continue;
            const bool linked = kind != 3 || __builtin_popcount(p ^ g) == 1;
            if (!linked) continue;
            NVLinkInfo l; l.LinkCount = 18; l.BandwidthGBps = bw; l.Version = 4; l.IsActive = lcg(rng) % 40 != 0;
            d.Topology.NVLinkConnections[uuid(idx, p)] = l;
        }
        d.Topology.NVLinkConnections["GPU-on-another-node"] = NVLinkInfo{1, 50.0, 4, true};   // ignored by the scorer
        d.Topology.PCIeInfo.BandwidthGBps = kind == 1 ? 32.0 : 64.0;
        d.Topology.PCIeInfo.SwitchID = switches[g / 2];
        d.Topology.NUMANode = numa_ids[g / 4];
        if (g == 3 && idx % 3 == 0) {
            d.MIGConfig = std::make_shared<MIGConfiguration>();
            d.MIGConfig->Enabled = true;
            d.MIGConfig->CurrentInstances = {{0, {"1g.10gb"}, false}, {1, {"1g.10gb"}, true}, {2, {"2g.20gb"}, false}, {3, {"weird"}, false}};
        }
        n.GPUs.push_back(d);
    }
    return n;
}

inline std::vector<NodeTopology> make_cluster(int n_nodes) {
    std::vector<NodeTopology> v; uint64_t rng = 12345;
    for (int i = 0; i < n_nodes; ++i) v.push_back(make_node(i, (int)(lcg(rng) % 4), rng));
    return v;
}

inline std::vector<GPUWorkload> make_workloads(int n) {
    std::vector<GPUWorkload> v; uint64_t rng = 777;
    const char* prefs[5] = {"None", "NVLinkOptimal", "NVLinkRequired", "SameNUMA", "SamePCIeSwitch"};
    for (int i = 0; i < n; ++i) {
        GPUWorkload w;
        char b[32]; snprintf(b, sizeof b, "job-%04d", i);
        w.Name = b; w.Namespace = "ml"; w.UID = std::string("uid-") + b;
        w.Requirements.GPUCount = 1 + (int)(lcg(rng) % 8);
        w.Requirements.MinMemoryGB = (int)(lcg(rng) % 3) * 30 + 10;
        w.Requirements.TopologyPreference = prefs[lcg(rng) % 5];
        if (lcg(rng) % 4 == 0) w.Requirements.GPUArchitecture = (lcg(rng) % 2) ? "Hopper" : "Blackwell";
        if (lcg(rng) % 10 == 0) { w.Requirements.MIGReq = std::make_shared<MIGRequirements>(); w.Requirements.MIGReq->Profile = "1g.10gb"; w.Requirements.MIGReq->Count = 1; }
        v.push_back(w);
    }
    return v;
}
}  // namespace fixture
