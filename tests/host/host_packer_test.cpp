// CPU-only: the packer (NodeTopology -> kgwe_node_image) and the string tables.
#include <cstdio>
#include <cstdlib>

#include "cluster_fixture.hpp"

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); exit(1); } } while (0)

int main() {
    using namespace kgwe_host;
    uint64_t rng = 1;
    NodeTopology n = fixture::make_node(3, 3, rng);          // cube-mesh A100, idx 3 -> GPU 3 carries MIG instances
    for (auto& g : n.GPUs) { g.HealthStatus = "Healthy"; for (auto& kv : g.Topology.NVLinkConnections) kv.second.IsActive = true; }
    n.GPUs[5].HealthStatus = "Unhealthy";
    n.GPUs[0].Topology.NVLinkConnections[fixture::uuid(3, 1)].IsActive = false;
    PackedNode pk;
    CHECK(PackNode(n, &pk).empty());
    const kgwe_node_image& im = pk.image;
    CHECK(im.n_gpus == 8 && pk.uuids.size() == 8 && pk.uuids[2] == fixture::uuid(3, 2));
    CHECK(im.healthy_mask == (0xFF & ~(1 << 5)));
    CHECK(im.nvl_present[0] == ((1 << 1) | (1 << 2) | (1 << 4)));          // cube edges of GPU 0; the off-node peer is dropped
    CHECK(im.nvl_active[0] == ((1 << 2) | (1 << 4)));
    CHECK(im.nvl_bw[0][1] == 600.0 && im.nvl_bw[0][3] == 0.0 && im.nvl_bw[6][7] == 600.0);
    CHECK(im.pcie_bw[4] == 64.0 && im.arch[0] == KGWE_ARCH_AMPERE);
    // NUMA {7,7,7,7,-1,-1,-1,-1} -> ranks {1,1,1,1,0,0,0,0}: order preserved, -1 first
    CHECK(im.numa[0] == 1 && im.numa[3] == 1 && im.numa[4] == 0 && im.numa[7] == 0);
    // switch ids in first-appearance order with "" -> "default": default=0, pcie-sw-B=1, pcie-sw-A=2, zz=3
    CHECK(im.pcie_switch[0] == 0 && im.pcie_switch[2] == 1 && im.pcie_switch[4] == 2 && im.pcie_switch[6] == 3);
    // free MIG instances on enabled GPUs: 1g.10gb x1 (the InUse one and the unknown name do not count), 2g.20gb x1
    CHECK(im.mig_free[1] == 1 && im.mig_free[3] == 1 && im.mig_free[0] == 0);
    CHECK(ArchCode("") == KGWE_ARCH_ANY && ArchCode("Blackwell") == KGWE_ARCH_BLACKWELL);
    const uint8_t rubin = ArchCode("Rubin");
    CHECK(rubin >= 7 && rubin == ArchCode("Rubin") && ArchCode("Feynman") == rubin + 1);
    CHECK(MIGProfileCode("7g.80gb") == 10 && MIGProfileCode("9g.1tb") == KGWE_MIG_UNKNOWN);
    CHECK(TopologyPreferenceCode("SameNUMA") == KGWE_TOPO_SAME_NUMA && TopologyPreferenceCode("bogus") == KGWE_TOPO_NONE);
    NodeTopology big = n; big.GPUs.push_back(n.GPUs[0]);
    CHECK(!PackNode(big, &pk).empty());
    // without a GPU the scheduler refuses to exist (no CPU fallback)
    auto r = TopologyAwareScheduler::New(SchedulerConfig{}, 16);
    if (!r.ok()) CHECK(r.error.find("no CPU fallback") != std::string::npos);
    printf("host_packer_test OK (engine %s)\n", r.ok() ? "created" : "refused: no GPU");
    return 0;
}
