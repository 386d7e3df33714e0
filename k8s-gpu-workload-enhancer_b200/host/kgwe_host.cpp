// kgwe_host.cpp — see kgwe_host.hpp.  Everything numeric happens behind the C ABI on the GPU;
// this file packs structs into the ABI's records, keeps the string tables, and rebuilds the
// reference's result types (names, UUIDs, reasons, error strings) from the ABI's indices.
#include "kgwe_host.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <set>

namespace kgwe_host {

namespace {
const char* kArch[] = {"", "Volta", "Turing", "Ampere", "Hopper", "Ada", "Blackwell"};
const char* kMig[KGWE_MIG_PROFILES] = {"1g.5gb", "1g.10gb", "2g.10gb", "2g.20gb", "3g.20gb", "3g.40gb", "4g.20gb", "4g.40gb", "7g.40gb", "7g.80gb"};
std::mutex g_arch_mu;
std::unordered_map<std::string, uint8_t> g_arch_extra;

std::vector<int> order_list(uint32_t order) {
    std::vector<int> v; const int n = (order >> 24) & 15;
    for (int p = 0; p < n; ++p) v.push_back((order >> (3 * p)) & 7);
    return v;
}
}  // namespace

uint8_t ArchCode(const std::string& arch) {
    for (int i = 0; i < 7; ++i) if (arch == kArch[i]) return (uint8_t)i;
    std::lock_guard<std::mutex> lk(g_arch_mu);
    auto it = g_arch_extra.find(arch);
    if (it != g_arch_extra.end()) return it->second;
    if (g_arch_extra.size() >= 247) return 255;
    const uint8_t code = (uint8_t)(7 + g_arch_extra.size());
    g_arch_extra[arch] = code;
    return code;
}

uint8_t MIGProfileCode(const std::string& name) {
    for (int i = 0; i < KGWE_MIG_PROFILES; ++i) if (name == kMig[i]) return (uint8_t)(i + 1);
    return KGWE_MIG_UNKNOWN;
}

uint8_t TopologyPreferenceCode(const std::string& p) {
    if (p == "NVLinkOptimal") return KGWE_TOPO_NVLINK_OPTIMAL;
    if (p == "NVLinkRequired") return KGWE_TOPO_NVLINK_REQUIRED;
    if (p == "SameNUMA") return KGWE_TOPO_SAME_NUMA;
    if (p == "SamePCIeSwitch") return KGWE_TOPO_SAME_PCIE_SWITCH;
    return KGWE_TOPO_NONE;          // "None" and anything else: the default arm, scheduler.go:325-331
}

std::string PackNode(const NodeTopology& node, PackedNode* out) {
    if (node.GPUs.size() > KGWE_MAX_GPUS) return "node " + node.NodeName + ": more than 8 GPUs";
    kgwe_node_image& im = out->image;
    memset(&im, 0, sizeof im);
    out->uuids.clear();
    std::unordered_map<std::string, int> gpu_of;
    for (size_t i = 0; i < node.GPUs.size(); ++i) { gpu_of[node.GPUs[i].UUID] = (int)i; out->uuids.push_back(node.GPUs[i].UUID); }
    // NUMA: order-preserving rank compression (only order and equality matter to scoreNUMATopology)
    std::set<int> numa_vals;
    for (auto& g : node.GPUs) numa_vals.insert(g.Topology.NUMANode);
    std::map<int, int> numa_rank; { int r = 0; for (int v : numa_vals) numa_rank[v] = r++; }
    // PCIe switch: interned in first-appearance order over node.GPUs, "" -> "default" (scheduler.go:484-486)
    std::unordered_map<std::string, uint8_t> sw;
    unsigned mig_free[KGWE_MIG_PROFILES] = {0};
    im.n_gpus = (uint8_t)node.GPUs.size();
    if (node.NVSwitchPresent) im.node_flags |= KGWE_NODE_NVSWITCH;
    for (size_t i = 0; i < node.GPUs.size(); ++i) {
        const GPUDevice& g = node.GPUs[i];
        for (auto& kv : g.Topology.NVLinkConnections) {
            auto it = gpu_of.find(kv.first);
            if (it == gpu_of.end() || it->second == (int)i) continue;    // peers outside the node never reach the scorer
            im.nvl_bw[i][it->second] = kv.second.BandwidthGBps;
            im.nvl_present[i] |= (uint8_t)(1u << it->second);
            if (kv.second.IsActive) im.nvl_active[i] |= (uint8_t)(1u << it->second);
        }
        im.pcie_bw[i] = g.Topology.PCIeInfo.BandwidthGBps;
        im.free_bytes[i] = g.MemoryFreeBytes;
        im.util[i] = g.UtilizationGPUPercent;
        std::string s = g.Topology.PCIeInfo.SwitchID.empty() ? "default" : g.Topology.PCIeInfo.SwitchID;
        if (!sw.count(s)) sw[s] = (uint8_t)sw.size();
        im.pcie_switch[i] = sw[s];
        im.numa[i] = (int8_t)numa_rank[g.Topology.NUMANode];
        im.arch[i] = ArchCode(g.Architecture);
        if (g.HealthStatus == "Healthy") im.healthy_mask |= (uint8_t)(1u << i);
        if (!g.Topology.NVLinkConnections.empty()) im.node_flags |= KGWE_NODE_ANY_NVLINK;
        if (g.MIGConfig && g.MIGConfig->Enabled) {
            im.mig_enabled_mask |= (uint8_t)(1u << i);
            for (auto& inst : g.MIGConfig->CurrentInstances) if (!inst.InUse) im.mig_has_free_mask |= (uint8_t)(1u << i);
        }
        if (g.MIGConfig && g.MIGConfig->Enabled)
            for (auto& inst : g.MIGConfig->CurrentInstances) {
                const uint8_t p = MIGProfileCode(inst.Profile.Name);
                if (!inst.InUse && p >= 1 && p <= KGWE_MIG_PROFILES) mig_free[p - 1]++;
            }
    }
    for (int p = 0; p < KGWE_MIG_PROFILES; ++p) {
        if (mig_free[p] > 255) return "node " + node.NodeName + ": more than 255 free MIG instances of one profile";
        im.mig_free[p] = (uint8_t)mig_free[p];
    }
    return "";
}

Result<std::unique_ptr<TopologyAwareScheduler>> TopologyAwareScheduler::New(const SchedulerConfig& cfg, uint32_t maxNodes, int device) {
    Result<std::unique_ptr<TopologyAwareScheduler>> r;
    kgwe_config c;
    kgwe_config_default(&c);
    c.device = device; c.max_nodes = maxNodes;
    c.topology_weight = cfg.TopologyWeight; c.resource_weight = cfg.ResourceWeight; c.balance_weight = cfg.BalanceWeight;
    kgwe_engine* e = nullptr;
    const int rc = kgwe_engine_create(&c, &e);
    if (rc != KGWE_OK) { r.error = "kgwe_engine_create failed (" + std::to_string(rc) + "): no usable CUDA device; the engine has no CPU fallback"; return r; }
    r.value.reset(new TopologyAwareScheduler());
    r.value->cfg_ = cfg; r.value->engine_ = e; r.value->max_nodes_ = maxNodes;
    return r;
}

TopologyAwareScheduler::~TopologyAwareScheduler() { if (engine_) kgwe_engine_destroy(engine_); }

uint32_t TopologyAwareScheduler::NodeIndex(const std::string& name) const {
    auto it = node_index_.find(name);
    return (it == node_index_.end() || node_names_[it->second].empty()) ? KGWE_NO_NODE : it->second;     // "" = currently removed
}

std::string TopologyAwareScheduler::UpsertNode(const NodeTopology& node) {
    PackedNode pk;
    std::string err = PackNode(node, &pk);
    if (!err.empty()) return err;
    auto known = node_index_.find(node.NodeName);       // a node that comes back keeps its index (and so its ledger entries, as in the reference)
    const bool fresh = known == node_index_.end();
    if (fresh && node_names_.size() >= max_nodes_) return "cluster image is full";
    const uint32_t idx = fresh ? (uint32_t)node_names_.size() : known->second;
    // the engine first: a record it rejects must not leave a name behind in the tables.  The reference keys allocations by
    // GPU UUID (scheduler.go:586-598), the engine by index in the node's GPU list: when a refresh changes that list (a GPU
    // fell off the bus, NVML enumerates in another order) the ledger entries of the node follow their GPUs.
    bool same_list = fresh || node_uuids_[idx].empty() || node_uuids_[idx] == pk.uuids;
    if (same_list) {
        if (kgwe_node_upsert(engine_, idx, &pk.image) != KGWE_OK) return kgwe_last_error(engine_);
    } else {
        int8_t old_to_new[KGWE_MAX_GPUS];
        for (int g = 0; g < KGWE_MAX_GPUS; ++g) {
            old_to_new[g] = -1;
            if (g >= (int)node_uuids_[idx].size()) continue;
            for (size_t q = 0; q < pk.uuids.size(); ++q) if (pk.uuids[q] == node_uuids_[idx][g]) old_to_new[g] = (int8_t)q;
        }
        if (kgwe_node_upsert_remap(engine_, idx, &pk.image, old_to_new) != KGWE_OK) return kgwe_last_error(engine_);
    }
    if (fresh) { node_index_[node.NodeName] = idx; node_names_.push_back(node.NodeName); node_uuids_.emplace_back(); }
    else node_names_[idx] = node.NodeName;
    node_uuids_[idx] = pk.uuids;
    return "";
}

std::string TopologyAwareScheduler::RemoveNode(const std::string& name) {
    const uint32_t idx = NodeIndex(name);
    if (idx == KGWE_NO_NODE) return "node " + name + " not found";
    if (kgwe_node_remove(engine_, idx) != KGWE_OK) return kgwe_last_error(engine_);
    node_names_[idx].clear();         // the name keeps its index for when the node comes back
    return "";
}

kgwe_pod_req TopologyAwareScheduler::PodReq(const GPUWorkload& w, bool commits) {
    kgwe_pod_req p; memset(&p, 0, sizeof p);
    // only calls that record an allocation need a ledger key; score-only calls (ScoreNodes, Filter,
    // Prioritize) must not grow the UID table with pods that may never be scheduled here
    auto it = wl_id_.find(w.UID);
    if (it == wl_id_.end() && commits) it = wl_id_.emplace(w.UID, next_wl_++).first;
    p.wl_id = it == wl_id_.end() ? 0 : it->second;
    p.gpu_count = w.Requirements.GPUCount; p.min_memory_gb = w.Requirements.MinMemoryGB;
    p.min_bandwidth_gbps = cfg_.EnableMinBandwidthFilter ? w.Requirements.MinBandwidthGBps : 0.0;
    p.topology_pref = TopologyPreferenceCode(w.Requirements.TopologyPreference);
    p.arch = w.Requirements.GPUArchitecture.empty() ? (uint8_t)KGWE_ARCH_ANY : ArchCode(w.Requirements.GPUArchitecture);
    if (w.Requirements.MIGReq) { p.mig_profile = MIGProfileCode(w.Requirements.MIGReq->Profile); p.mig_count = w.Requirements.MIGReq->Count; }
    return p;
}

std::vector<uint8_t> TopologyAwareScheduler::HintBitmap(const PlacementHint* hint) const {
    std::vector<uint8_t> bm;
    if (!hint) return bm;
    bm.assign((node_names_.size() + 7) / 8, 0);
    for (auto& n : hint->PreferredNodes) { const uint32_t i = NodeIndex(n); if (i != KGWE_NO_NODE) bm[i / 8] |= (uint8_t)(1u << (i & 7)); }
    return bm;
}

NodeScore TopologyAwareScheduler::NodeScoreOf(const kgwe_placement& pl) const {
    NodeScore s;
    s.NodeName = node_names_[pl.node_idx];
    s.TotalScore = pl.total; s.TopologyScore = pl.topology; s.ResourceScore = pl.resource; s.BalanceScore = pl.balance;
    for (int g : order_list(pl.order)) s.SelectedGPUs.push_back(node_uuids_[pl.node_idx][g]);
    // scheduler.go:270-297
    if (pl.flags & KGWE_F_HINT) s.Reasons.push_back("ML optimizer preferred this node");
    if (pl.topology >= 80) s.Reasons.push_back("Optimal NVLink connectivity available");
    if (pl.resource >= 80) s.Reasons.push_back("Sufficient resources with headroom");
    if (pl.balance >= 80) s.Reasons.push_back("Good load distribution");
    return s;
}

SchedulingDecision TopologyAwareScheduler::DecisionOf(const GPUWorkload& w, const kgwe_placement& pl) const {
    const NodeScore s = NodeScoreOf(pl);
    SchedulingDecision d;
    d.WorkloadUID = w.UID; d.SelectedNode = s.NodeName; d.SelectedGPUs = s.SelectedGPUs; d.Score = s.TotalScore;
    d.Reasons = s.Reasons; d.EstimatedBandwidth = pl.est_bandwidth_gbps;
    return d;
}

void TopologyAwareScheduler::UpdateLatency(double ms) {   // scheduler.go:805-819 ("P99" is the max there too)
    const double total = (double)(metrics_.TotalScheduled + metrics_.TotalFailed);
    metrics_.AverageLatencyMs = total == 0 ? ms : (metrics_.AverageLatencyMs * total + ms) / (total + 1);
    metrics_.P99LatencyMs = std::max(metrics_.P99LatencyMs, ms);
}

Result<std::vector<NodeScore>> TopologyAwareScheduler::ScoreNodes(const GPUWorkload& w, const PlacementHint* hint, uint32_t topk) {
    Result<std::vector<NodeScore>> r;
    if (!HasTopology()) { r.error = "cluster topology not available"; return r; }
    topk = std::min<uint32_t>(topk, KGWE_MAX_TOPK);
    const kgwe_pod_req p = PodReq(w, false);
    std::vector<uint8_t> bm = HintBitmap(hint);
    std::vector<kgwe_placement> out(topk);
    if (kgwe_score_batch(engine_, &p, 1, hint ? bm.data() : nullptr, nullptr, topk, out.data()) != KGWE_OK) { r.error = kgwe_last_error(engine_); return r; }
    for (auto& pl : out) { if (pl.node_idx == KGWE_NO_NODE) break; r.value.push_back(NodeScoreOf(pl)); }
    return r;
}

Result<SchedulingDecision> TopologyAwareScheduler::Schedule(const GPUWorkload& w, const PlacementHint* hint) {
    const auto t0 = std::chrono::steady_clock::now();
    Result<SchedulingDecision> r;
    if (!HasTopology()) { r.error = "cluster topology not available"; return r; }            // scheduler.go:124
    const bool known_uid = wl_id_.count(w.UID) != 0;
    const kgwe_pod_req p = PodReq(w, true);
    std::vector<uint8_t> bm = HintBitmap(hint);
    kgwe_placement pl;
    const int rc = kgwe_schedule_queue(engine_, &p, 1, hint ? bm.data() : nullptr, &pl);
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (rc != KGWE_OK) { r.error = kgwe_last_error(engine_); return r; }
    if (pl.node_idx == KGWE_NO_NODE) {
        UpdateLatency(ms); metrics_.TotalFailed++;
        if (!known_uid) wl_id_.erase(w.UID);                                                   // nothing was recorded under it
        r.error = "no eligible nodes found for workload " + w.Name;                            // scheduler.go:140
        return r;
    }
    r.value = DecisionOf(w, pl);
    UpdateLatency(ms); metrics_.TotalScheduled++;
    if (pl.topology >= 80) metrics_.TopologyOptimalPlacements++;                               // scheduler.go:154
    events_.push_back(SchedulingEvent{"Scheduled", w.UID, std::make_shared<SchedulingDecision>(r.value), ""});
    return r;
}

std::vector<Result<SchedulingDecision>> TopologyAwareScheduler::ScheduleBurst(const std::vector<GPUWorkload>& ws) {
    std::vector<Result<SchedulingDecision>> out(ws.size());
    if (ws.empty()) return out;
    if (!HasTopology()) { for (auto& r : out) r.error = "cluster topology not available"; return out; }
    std::vector<kgwe_pod_req> pods; pods.reserve(ws.size());
    std::set<std::string> fresh_uids;                       // UIDs this burst brought in; forgotten again if none of their tries lands
    for (auto& w : ws) { if (!wl_id_.count(w.UID)) fresh_uids.insert(w.UID); pods.push_back(PodReq(w, true)); }
    std::vector<kgwe_placement> pls(ws.size());
    if (kgwe_schedule_queue(engine_, pods.data(), (uint32_t)pods.size(), nullptr, pls.data()) != KGWE_OK) {
        for (auto& r : out) r.error = kgwe_last_error(engine_);
        return out;
    }
    for (size_t i = 0; i < ws.size(); ++i) {
        if (pls[i].node_idx == KGWE_NO_NODE) { metrics_.TotalFailed++; out[i].error = "no eligible nodes found for workload " + ws[i].Name; continue; }
        out[i].value = DecisionOf(ws[i], pls[i]);
        metrics_.TotalScheduled++;
        if (pls[i].topology >= 80) metrics_.TopologyOptimalPlacements++;
        events_.push_back(SchedulingEvent{"Scheduled", ws[i].UID, std::make_shared<SchedulingDecision>(out[i].value), ""});
        fresh_uids.erase(ws[i].UID);
    }
    for (auto& uid : fresh_uids) wl_id_.erase(uid);
    return out;
}

Result<TopologyHint> TopologyAwareScheduler::GetTopologyHint(const HintRequirements& req) {
    Result<TopologyHint> r;
    kgwe_hint_req q; memset(&q, 0, sizeof q);
    q.count = req.Count; q.min_memory_gb = req.MinMemoryGB; q.prefer_nvlink = req.PreferNVLink; q.prefer_same_numa = req.PreferSameNUMA;
    q.arch = req.Architecture.empty() ? (uint8_t)KGWE_ARCH_ANY : ArchCode(req.Architecture);
    kgwe_topology_hint t;
    if (kgwe_topology_hint_batch(engine_, &q, 1, &t) != KGWE_OK) { r.error = kgwe_last_error(engine_); return r; }
    if (t.node_idx == KGWE_NO_NODE) { r.error = "no suitable placement found for requirements"; return r; }     // discovery.go:241
    r.value.PreferredNode = node_names_[t.node_idx]; r.value.Score = t.score; r.value.BandwidthEstimate = t.bandwidth_estimate;
    for (int g : order_list(t.order)) r.value.PreferredGPUs.push_back(node_uuids_[t.node_idx][g]);
    // explainPlacement, discovery.go:544-558
    if (t.score >= 80) r.value.Reasons.push_back("Optimal NVLink connectivity between selected GPUs");
    else if (t.score >= 60) r.value.Reasons.push_back("Good topology placement with partial NVLink connectivity");
    else r.value.Reasons.push_back("PCIe-based connectivity (NVLink not available)");
    if (t.bandwidth_estimate > 500) { char b[64]; snprintf(b, sizeof b, "High aggregate bandwidth: %.0f GB/s", t.bandwidth_estimate); r.value.Reasons.push_back(b); }
    return r;
}

Result<SchedulingDecision> TopologyAwareScheduler::Bind(const GPUWorkload& w, const std::string& nodeName) {
    const auto t0 = std::chrono::steady_clock::now();
    Result<SchedulingDecision> r;
    if (!HasTopology()) { r.error = "cluster topology not available"; return r; }
    const uint32_t idx = NodeIndex(nodeName);
    if (idx == KGWE_NO_NODE) { r.error = "node " + nodeName + " is not in the GPU topology snapshot"; return r; }
    kgwe_stats st;
    if (kgwe_get_stats(engine_, &st) != KGWE_OK) { r.error = kgwe_last_error(engine_); return r; }
    std::vector<kgwe_node_score> sc(st.n_nodes);
    std::vector<kgwe_placement> top(KGWE_MAX_TOPK);
    const bool known_uid = wl_id_.count(w.UID) != 0;
    const kgwe_pod_req p = PodReq(w, true);
    auto forget = [&] { if (!known_uid) wl_id_.erase(w.UID); };
    if (kgwe_score_batch(engine_, &p, 1, nullptr, sc.data(), KGWE_MAX_TOPK, top.data()) != KGWE_OK) { forget(); r.error = kgwe_last_error(engine_); return r; }
    if (idx >= sc.size() || !(sc[idx].flags & KGWE_F_KEPT)) {
        forget(); metrics_.TotalFailed++;
        r.error = "failed to schedule workload " + w.Name + " after trying all nodes";          // scheduler.go:178 (the one node it was asked to try)
        return r;
    }
    if (kgwe_alloc_commit(engine_, p.wl_id, idx, sc[idx].gpu_mask) != KGWE_OK) { forget(); metrics_.TotalFailed++; r.error = kgwe_last_error(engine_); return r; }
    kgwe_placement pl; memset(&pl, 0, sizeof pl);
    pl.node_idx = idx; pl.total = sc[idx].total; pl.topology = sc[idx].topology; pl.resource = sc[idx].resource; pl.balance = sc[idx].balance;
    pl.gpu_mask = sc[idx].gpu_mask; pl.flags = sc[idx].flags; pl.order = sc[idx].order;
    for (auto& t : top) if (t.node_idx == idx) pl.est_bandwidth_gbps = t.est_bandwidth_gbps;    // (known when the node is among the best 16)
    r.value = DecisionOf(w, pl);
    UpdateLatency(std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    metrics_.TotalScheduled++;
    if (pl.topology >= 80) metrics_.TopologyOptimalPlacements++;
    events_.push_back(SchedulingEvent{"Scheduled", w.UID, std::make_shared<SchedulingDecision>(r.value), ""});
    return r;
}

std::string TopologyAwareScheduler::MetricsText() {
    kgwe_stats st;
    if (kgwe_get_stats(engine_, &st) != KGWE_OK) return "# kgwe_get_stats failed\n";
    static const char* edges[KGWE_LATENCY_BUCKETS] = {"10", "25", "50", "100", "250", "500", "1000", "2500", "5000", "+Inf"};
    std::string o;
    char b[256];
    o += "# HELP kgwe_scheduling_latency_us Histogram of engine call latency in microseconds (the reference's kgwe_scheduling_latency_ms buckets, scaled)\n# TYPE kgwe_scheduling_latency_us histogram\n";
    uint64_t cum = 0;
    for (int i = 0; i < KGWE_LATENCY_BUCKETS; ++i) { cum += st.latency_us_bucket[i]; snprintf(b, sizeof b, "kgwe_scheduling_latency_us_bucket{le=\"%s\"} %llu\n", edges[i], (unsigned long long)cum); o += b; }
    snprintf(b, sizeof b, "kgwe_scheduling_latency_us_sum %.3f\nkgwe_scheduling_latency_us_count %llu\n", st.latency_sum_us, (unsigned long long)st.latency_count); o += b;
    auto counter = [&](const char* name, const char* help, long long v) { snprintf(b, sizeof b, "# HELP %s %s\n# TYPE %s counter\n%s %lld\n", name, help, name, name, v); o += b; };
    counter("kgwe_scheduling_attempts_total", "Total number of scheduling attempts", (long long)(metrics_.TotalScheduled + metrics_.TotalFailed));
    counter("kgwe_scheduling_successes_total", "Total number of successful schedulings", (long long)metrics_.TotalScheduled);
    counter("kgwe_scheduling_failures_total", "Total number of scheduling failures", (long long)metrics_.TotalFailed);
    counter("kgwe_topology_optimal_placements_total", "Placements with TopologyScore >= 80", (long long)metrics_.TopologyOptimalPlacements);
    counter("kgwe_engine_evaluations_total", "(pod, node) pairs scored by the engine", (long long)st.evaluations);
    counter("kgwe_engine_candidates_total", "GPU subsets whose score the engine computed", (long long)st.candidates);
    return o;
}

Result<std::map<std::string, int>> TopologyAwareScheduler::TopologyScores() {
    Result<std::map<std::string, int>> r;
    std::vector<int32_t> v(node_names_.size());
    if (!v.empty() && kgwe_topology_gauges(engine_, v.data(), (uint32_t)v.size()) != KGWE_OK) { r.error = kgwe_last_error(engine_); return r; }
    for (size_t i = 0; i < v.size(); ++i) if (!node_names_[i].empty()) r.value[node_names_[i]] = v[i];
    return r;
}

std::string TopologyAwareScheduler::ScoreAll(const GPUWorkload& w, std::vector<kgwe_node_score>* scores) {
    if (!HasTopology()) return "cluster topology not available";
    kgwe_stats st;
    if (kgwe_get_stats(engine_, &st) != KGWE_OK) return kgwe_last_error(engine_);
    scores->assign(st.n_nodes, kgwe_node_score{});
    const kgwe_pod_req p = PodReq(w, false);
    if (st.n_nodes && kgwe_score_batch(engine_, &p, 1, nullptr, scores->data(), 0, nullptr) != KGWE_OK) return kgwe_last_error(engine_);
    return "";
}

ExtenderFilterResult TopologyAwareScheduler::Filter(const GPUWorkload& w, const std::vector<std::string>& names) {
    ExtenderFilterResult r;
    std::vector<kgwe_node_score> sc;
    r.Error = ScoreAll(w, &sc);
    if (!r.Error.empty()) return r;
    for (auto& n : names) {
        const uint32_t i = NodeIndex(n);
        if (i == KGWE_NO_NODE || i >= sc.size()) { r.FailedNodes[n] = "node not in the GPU topology snapshot"; continue; }
        if (sc[i].flags & KGWE_F_KEPT) r.NodeNames.push_back(n);
        else if (sc[i].flags & KGWE_F_MINBW) r.FailedNodes[n] = "inter-GPU bandwidth below the requested minimum";
        else if (sc[i].flags & KGWE_F_ELIGIBLE) r.FailedNodes[n] = "placement score is not positive";
        else r.FailedNodes[n] = "not enough available GPUs of the required kind";
    }
    return r;
}

std::vector<HostPriority> TopologyAwareScheduler::Prioritize(const GPUWorkload& w, const std::vector<std::string>& names) {
    std::vector<HostPriority> out;
    std::vector<kgwe_node_score> sc;
    const std::string err = ScoreAll(w, &sc);
    for (auto& n : names) {
        HostPriority hp; hp.Host = n;
        const uint32_t i = NodeIndex(n);
        if (err.empty() && i != KGWE_NO_NODE && i < sc.size() && (sc[i].flags & KGWE_F_KEPT)) {
            const int64_t t = std::min<int64_t>(100, std::max<int64_t>(0, sc[i].total));
            hp.Score = t * MaxExtenderPriority / 100;
        }
        out.push_back(hp);
    }
    return out;
}

std::string TopologyAwareScheduler::ReleaseAllocation(const std::string& uid) {
    auto it = wl_id_.find(uid);
    if (it == wl_id_.end() || kgwe_alloc_release(engine_, it->second) != KGWE_OK)
        return "allocation not found for workload " + uid;                                      // scheduler.go:715
    wl_id_.erase(it);
    events_.push_back(SchedulingEvent{"Released", uid, nullptr, ""});
    return "";
}

}  // namespace kgwe_host
