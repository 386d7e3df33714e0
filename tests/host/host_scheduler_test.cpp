// GPU: the C++ mirror of TopologyAwareScheduler (Schedule / ScoreNodes / ReleaseAllocation / metrics /
// events) on a struct-form cluster, checked against the CPU oracle run on the packed images.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>

#include "../../oracle/kgwe_oracle.h"
#include "cluster_fixture.hpp"

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); exit(1); } } while (0)

using namespace kgwe_host;

static kgwe_pod_req pod_of(const GPUWorkload& w, uint64_t id) {
    kgwe_pod_req p; memset(&p, 0, sizeof p);
    p.wl_id = id; p.gpu_count = w.Requirements.GPUCount; p.min_memory_gb = w.Requirements.MinMemoryGB;
    p.topology_pref = TopologyPreferenceCode(w.Requirements.TopologyPreference);
    p.arch = w.Requirements.GPUArchitecture.empty() ? (uint8_t)KGWE_ARCH_ANY : ArchCode(w.Requirements.GPUArchitecture);
    if (w.Requirements.MIGReq) { p.mig_profile = MIGProfileCode(w.Requirements.MIGReq->Profile); p.mig_count = w.Requirements.MIGReq->Count; }
    return p;
}

int main() {
    const int N = 48, P = 160;
    auto cluster = fixture::make_cluster(N);
    auto jobs = fixture::make_workloads(P);
    auto made = TopologyAwareScheduler::New(SchedulerConfig{}, 64);
    CHECK(made.ok());
    auto& S = *made.value;
    GPUWorkload early = jobs[0];
    CHECK(S.Schedule(early).error == "cluster topology not available");                 // scheduler.go:124
    std::vector<kgwe_node_image> images;
    for (auto& n : cluster) { CHECK(S.UpsertNode(n).empty()); PackedNode pk; CHECK(PackNode(n, &pk).empty()); images.push_back(pk.image); }

    // oracle on the same packed images, same order of workloads
    std::vector<kgwe_pod_req> pods;
    for (int i = 0; i < P; ++i) pods.push_back(pod_of(jobs[i], (uint64_t)i + 1));
    std::vector<kgwe_placement> want(P);
    std::vector<kgwe_oracle_alloc> led(P); uint32_t n_led = 0;
    CHECK(kgwe_oracle_schedule_queue(nullptr, images.data(), N, nullptr, 0, pods.data(), P, nullptr, want.data(), led.data(), &n_led, KGWE_ORACLE_FAITHFUL, nullptr) == 0);

    // Prioritize before anything is placed: ranked NodeScores for job 5 against the oracle's top-4
    {
        std::vector<kgwe_placement> top(4);
        CHECK(kgwe_oracle_score_batch(nullptr, images.data(), N, nullptr, 0, &pods[5], 1, nullptr, nullptr, 4, top.data(), KGWE_ORACLE_INDEXED, 1, nullptr) == 0);
        auto ranked = S.ScoreNodes(jobs[5], nullptr, 4);
        CHECK(ranked.ok());
        size_t n_valid = 0; for (auto& t : top) if (t.node_idx != KGWE_NO_NODE) ++n_valid;
        CHECK(ranked.value.size() == n_valid);
        for (size_t i = 0; i < n_valid; ++i) {
            CHECK(ranked.value[i].NodeName == cluster[top[i].node_idx].NodeName && ranked.value[i].TotalScore == top[i].total);
            CHECK(ranked.value[i].TopologyScore == top[i].topology && ranked.value[i].BalanceScore == top[i].balance);
        }
    }
    // first half one Schedule() call per workload, second half as one burst
    int placed = 0, failed = 0, optimal = 0;
    std::vector<Result<SchedulingDecision>> got;
    for (int i = 0; i < P / 2; ++i) got.push_back(S.Schedule(jobs[i]));
    { std::vector<GPUWorkload> rest(jobs.begin() + P / 2, jobs.end()); for (auto& r : S.ScheduleBurst(rest)) got.push_back(r); }
    for (int i = 0; i < P; ++i) {
        if (want[i].node_idx == KGWE_NO_NODE) {
            CHECK(!got[i].ok() && got[i].error == "no eligible nodes found for workload " + jobs[i].Name);   // scheduler.go:140
            ++failed; continue;
        }
        CHECK(got[i].ok());
        const SchedulingDecision& d = got[i].value;
        CHECK(d.WorkloadUID == jobs[i].UID && d.SelectedNode == cluster[want[i].node_idx].NodeName && d.Score == want[i].total);
        const int n = (want[i].order >> 24) & 15;
        CHECK((int)d.SelectedGPUs.size() == n);
        for (int q = 0; q < n; ++q) CHECK(d.SelectedGPUs[q] == fixture::uuid((int)want[i].node_idx, (want[i].order >> (3 * q)) & 7));
        CHECK(d.EstimatedBandwidth == want[i].est_bandwidth_gbps);
        const bool top80 = want[i].topology >= 80;
        bool has = false; for (auto& r : d.Reasons) if (r == "Optimal NVLink connectivity available") has = true;
        CHECK(has == top80);
        ++placed; if (top80) ++optimal;
    }
    CHECK(placed > 20 && failed > 0);
    SchedulerMetrics m = S.GetMetrics();
    CHECK(m.TotalScheduled == placed && m.TotalFailed == failed && m.TopologyOptimalPlacements == optimal);
    CHECK((int)S.Events().size() == placed && S.Events().front().Type == "Scheduled");
    // release: the GPUs come back; unknown / repeated release -> the reference's error string
    int first_placed = 0; while (want[first_placed].node_idx == KGWE_NO_NODE) ++first_placed;
    CHECK(S.ReleaseAllocation(jobs[first_placed].UID).empty());
    CHECK(S.ReleaseAllocation(jobs[first_placed].UID) == "allocation not found for workload " + jobs[first_placed].UID);   // scheduler.go:715
    CHECK(S.ReleaseAllocation("never-seen") == "allocation not found for workload never-seen");
    CHECK(S.Events().back().Type == "Released");
    // a hint turns into +10 and a reason (scheduler.go:270-279)
    {
        GPUWorkload w = jobs[1]; w.UID = "uid-hinted"; w.Name = "hinted"; w.Requirements.GPUCount = 1; w.Requirements.TopologyPreference = "None";
        w.Requirements.GPUArchitecture.clear(); w.Requirements.MIGReq.reset();
        auto plain = S.ScoreNodes(w, nullptr, 16);
        CHECK(plain.ok() && plain.value.size() >= 2);
        PlacementHint h; h.PreferredNodes = {plain.value.back().NodeName, "no-such-node"};
        auto hinted = S.ScoreNodes(w, &h, 16);
        CHECK(hinted.ok());
        bool seen = false;
        for (auto& s : hinted.value) if (s.NodeName == plain.value.back().NodeName) { seen = true; CHECK(s.TotalScore == plain.value.back().TotalScore + 10); CHECK(s.Reasons[0] == "ML optimizer preferred this node"); }
        CHECK(seen);
    }
    // GetTopologyHint / exporter gauge through the mirror, against the oracle on the packed images
    {
        HintRequirements hr; hr.Count = 4; hr.MinMemoryGB = 40; hr.PreferNVLink = true; hr.PreferSameNUMA = true; hr.Architecture = "Blackwell";
        kgwe_hint_req q; memset(&q, 0, sizeof q);
        q.count = 4; q.min_memory_gb = 40; q.prefer_nvlink = 1; q.prefer_same_numa = 1; q.arch = KGWE_ARCH_BLACKWELL;
        kgwe_topology_hint t;
        CHECK(kgwe_oracle_topology_hints(nullptr, images.data(), N, &q, 1, &t) == 0);
        auto h = S.GetTopologyHint(hr);
        CHECK(t.node_idx != KGWE_NO_NODE && h.ok());
        CHECK(h.value.PreferredNode == cluster[t.node_idx].NodeName && h.value.Score == t.score && h.value.BandwidthEstimate == t.bandwidth_estimate);
        CHECK((int)h.value.PreferredGPUs.size() == 4 && h.value.PreferredGPUs[0] == fixture::uuid((int)t.node_idx, t.order & 7));
        CHECK(h.value.Reasons[0] == (t.score >= 80 ? "Optimal NVLink connectivity between selected GPUs" : "Good topology placement with partial NVLink connectivity"));
        hr.Count = 9;
        CHECK(S.GetTopologyHint(hr).error == "no suitable placement found for requirements");
        std::vector<int32_t> gauges(N);
        CHECK(kgwe_oracle_topology_gauges(images.data(), N, gauges.data()) == 0);
        auto ts = S.TopologyScores();
        CHECK(ts.ok() && (int)ts.value.size() == N);
        for (int i = 0; i < N; ++i) CHECK(ts.value[cluster[i].NodeName] == gauges[i]);
    }
    // extender verbs: payloads built from the score matrix of one pod
    {
        GPUWorkload w = jobs[7]; w.UID = "uid-ext"; w.Name = "ext";
        kgwe_pod_req q = pod_of(w, 999999);
        std::vector<kgwe_node_score> want_sc(N);
        // current ledger = oracle's final ledger minus the released workload
        std::vector<kgwe_oracle_alloc> cur;
        for (uint32_t i = 0; i < n_led; ++i) if (led[i].wl_id != (uint64_t)first_placed + 1) cur.push_back(led[i]);
        CHECK(kgwe_oracle_score_batch(nullptr, images.data(), N, cur.data(), (uint32_t)cur.size(), &q, 1, nullptr, want_sc.data(), 0, nullptr, KGWE_ORACLE_INDEXED, 1, nullptr) == 0);
        std::vector<std::string> names;
        for (int i = 0; i < N; i += 2) names.push_back(cluster[i].NodeName);
        names.push_back("ghost-node");
        ExtenderFilterResult fr = S.Filter(w, names);
        CHECK(fr.Error.empty() && fr.FailedNodes.count("ghost-node") == 1);
        size_t kept = 0;
        for (int i = 0; i < N; i += 2) {
            const bool k = want_sc[i].flags & KGWE_F_KEPT;
            kept += k;
            CHECK((std::find(fr.NodeNames.begin(), fr.NodeNames.end(), cluster[i].NodeName) != fr.NodeNames.end()) == k);
            CHECK((fr.FailedNodes.count(cluster[i].NodeName) == 1) == !k);
        }
        CHECK(fr.NodeNames.size() == kept);
        auto pr = S.Prioritize(w, names);
        CHECK(pr.size() == names.size() && pr.back().Host == "ghost-node" && pr.back().Score == 0);
        for (size_t x = 0; x + 1 < pr.size(); ++x) {
            const kgwe_node_score& ws = want_sc[2 * x];
            const int64_t exp = (ws.flags & KGWE_F_KEPT) ? std::min<int64_t>(100, std::max<int64_t>(0, ws.total)) * 10 / 100 : 0;
            CHECK(pr[x].Host == cluster[2 * x].NodeName && pr[x].Score == exp && pr[x].Score <= MaxExtenderPriority);
        }
    }
    // a record the engine rejects leaves no name behind, and the next new node takes the index it would have had
    {
        NodeTopology bad = cluster[1]; bad.NodeName = "bad-node";
        bad.GPUs[0].Topology.PCIeInfo.BandwidthGBps = -5.0;
        CHECK(!S.UpsertNode(bad).empty() && S.NodeIndex("bad-node") == KGWE_NO_NODE);
        NodeTopology extra = cluster[1]; extra.NodeName = "extra-node";
        CHECK(S.UpsertNode(extra).empty() && S.NodeIndex("extra-node") == (uint32_t)N);
        CHECK(S.RemoveNode("extra-node").empty());
    }
    // a released workload can be scheduled again under the same UID (fresh ledger key), and released again
    {
        auto again = S.Schedule(jobs[first_placed]);
        CHECK(again.ok() && again.value.WorkloadUID == jobs[first_placed].UID);
        CHECK(S.ReleaseAllocation(jobs[first_placed].UID).empty());
        CHECK(S.ReleaseAllocation(jobs[first_placed].UID) == "allocation not found for workload " + jobs[first_placed].UID);
    }
    // a refresh that re-orders a node's GPU list: allocations are keyed by UUID in the reference (scheduler.go:586-598),
    // so the GPUs a workload holds stay taken wherever they now sit in the list
    {
        auto made2 = TopologyAwareScheduler::New(SchedulerConfig{}, 4);
        CHECK(made2.ok());
        auto& T = *made2.value;
        uint64_t rng = 99;
        NodeTopology n = fixture::make_node(0, 0, rng);
        for (auto& g : n.GPUs) { g.HealthStatus = "Healthy"; for (auto& kv : g.Topology.NVLinkConnections) kv.second.IsActive = true; }
        CHECK(T.UpsertNode(n).empty());
        GPUWorkload w; w.Name = "holder"; w.UID = "uid-holder"; w.Requirements.GPUCount = 3; w.Requirements.MinMemoryGB = 1; w.Requirements.TopologyPreference = "NVLinkOptimal";
        auto d = T.Schedule(w);
        CHECK(d.ok() && d.value.SelectedGPUs.size() == 3);
        const std::vector<std::string> held = d.value.SelectedGPUs;
        NodeTopology r = n;
        std::reverse(r.GPUs.begin(), r.GPUs.end());                      // NVML enumerates the other way round after a driver restart
        r.GPUs.erase(r.GPUs.begin() + 7);                                 // ... and the GPU formerly at index 0 is gone
        for (size_t g = 0; g < r.GPUs.size(); ++g) r.GPUs[g].Index = (int)g;
        CHECK(T.UpsertNode(r).empty());
        GPUWorkload all; all.Name = "rest"; all.UID = "uid-rest"; all.Requirements.GPUCount = 5; all.Requirements.MinMemoryGB = 1; all.Requirements.TopologyPreference = "None";
        auto d2 = T.Schedule(all);
        // 7 GPUs left, of which the holder still has those of its three that survived
        int survived = 0; for (auto& u : held) if (u != fixture::uuid(0, 0)) ++survived;
        if (7 - survived >= 5) {
            CHECK(d2.ok());
            for (auto& u : d2.value.SelectedGPUs) CHECK(std::find(held.begin(), held.end(), u) == held.end());
        } else CHECK(!d2.ok());
        CHECK(T.ReleaseAllocation("uid-holder").empty());
        GPUWorkload big; big.Name = "big"; big.UID = "uid-big"; big.Requirements.GPUCount = 7 - (d2.ok() ? 5 : 0); big.Requirements.MinMemoryGB = 1; big.Requirements.TopologyPreference = "None";
        CHECK(T.Schedule(big).ok());                                       // everything the holder had is free again
    }
    // node delete event
    CHECK(S.RemoveNode(cluster[0].NodeName).empty() && S.NodeIndex(cluster[0].NodeName) == KGWE_NO_NODE);
    CHECK(!S.RemoveNode(cluster[0].NodeName).empty());
    printf("host_scheduler_test OK: %d placed, %d failed, %d topology-optimal, %zu events\n", placed, failed, optimal, S.Events().size());
    return 0;
}
