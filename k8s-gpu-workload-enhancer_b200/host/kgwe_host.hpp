// kgwe_host.hpp — C++ mirror of the reference's scheduler interface for the scoring path,
// sitting above the C ABI (include/kgwe_engine.h).  The reference's host code is Go and no Go
// toolchain exists in this image, so this is the compiled host side: same type and method
// names, argument meaning and error strings as
//   src/scheduler/scheduler.go  (TopologyAwareScheduler: Schedule :114, ReleaseAllocation :710,
//                                GetMetrics :793, Events :800)
//   src/scheduler/types.go      (GPUWorkload :13, GPURequirements :36, SchedulingDecision :253,
//                                NodeScore :295, SchedulerMetrics :322, SchedulerConfig :346)
//   src/discovery/types.go      (GPUDevice :11, DeviceTopology :116, NVLinkInfo :134,
//                                PCIeTopology :149, MIGConfiguration :167, NodeTopology :351)
// It also contains the packer (GPUDevice structs -> kgwe_node_image) and the
// string <-> index tables; strings never cross the C ABI.
#pragma once
#include <cstdint>
#include <deque>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/kgwe_engine.h"

namespace kgwe_host {

// ---- discovery types (only the fields the scorer reads) ----
struct NVLinkInfo { int LinkCount = 0; double BandwidthGBps = 0; int Version = 0; bool IsActive = false; };
struct PCIeTopology { std::string BusID; int Generation = 0; int Width = 0; double BandwidthGBps = 0; std::string SwitchID; };
struct MIGProfile { std::string Name; };
struct MIGInstance { int InstanceID = 0; MIGProfile Profile; bool InUse = false; };
struct MIGConfiguration { bool Enabled = false; std::vector<MIGInstance> CurrentInstances; };
struct DeviceTopology { std::map<std::string, NVLinkInfo> NVLinkConnections; PCIeTopology PCIeInfo; int NUMANode = -1; };
struct GPUDevice {
    std::string UUID; int Index = 0; std::string Name; std::string Architecture;
    uint64_t MemoryFreeBytes = 0;            // Memory.FreeBytes
    double UtilizationGPUPercent = 0;        // Utilization.GPUPercent
    std::string HealthStatus = "Healthy";    // Health.Status
    DeviceTopology Topology;
    std::shared_ptr<MIGConfiguration> MIGConfig;   // nil-able, like the Go pointer
};
struct NodeTopology { std::string NodeName; std::vector<GPUDevice> GPUs; bool NVSwitchPresent = false; /* NVSwitchInfo != nil && Present */ };

// ---- scheduler types ----
struct MIGRequirements { std::string Profile; int Count = 0; };
struct GPURequirements {
    int GPUCount = 0; int MinMemoryGB = 0; double MinBandwidthGBps = 0;
    std::string TopologyPreference = "None";
    std::shared_ptr<MIGRequirements> MIGReq;   // MIGRequirements *MIGRequirements
    std::string GPUArchitecture;
};
struct GPUWorkload { std::string Name, Namespace, UID; GPURequirements Requirements; std::string WorkloadType = "Training"; int32_t Priority = 0; };
struct PlacementHint { std::vector<std::string> PreferredNodes; };
struct SchedulingDecision {
    std::string WorkloadUID, SelectedNode; std::vector<std::string> SelectedGPUs;
    int Score = 0; std::vector<std::string> Reasons; double EstimatedBandwidth = 0;
};
struct NodeScore {
    std::string NodeName; int TotalScore = 0, TopologyScore = 0, ResourceScore = 0, LocalityScore = 0, BalanceScore = 0;
    std::vector<std::string> SelectedGPUs; std::vector<std::string> Reasons;
};
struct SchedulerMetrics {
    int64_t TotalScheduled = 0, TotalFailed = 0; double AverageLatencyMs = 0, P99LatencyMs = 0;
    int64_t TopologyOptimalPlacements = 0, MIGAllocations = 0, PreemptionCount = 0;
};
struct SchedulerConfig {     // DefaultSchedulerConfig(), types.go:379-392
    std::string SchedulerName = "kgwe-scheduler";
    int TopologyWeight = 40, ResourceWeight = 35, BalanceWeight = 25;
    bool EnableMinBandwidthFilter = false;   // extension E3: honour Requirements.MinBandwidthGBps
};
struct SchedulingEvent { std::string Type, WorkloadUID; std::shared_ptr<SchedulingDecision> Decision; std::string Error; };

// discovery.GPURequirements / TopologyHint (src/discovery/discovery.go:250-281, src/discovery/types.go TopologyHint):
// the request and answer of DiscoveryService.GetTopologyHint (discovery.go:222-247).
struct HintRequirements { int Count = 0; int MinMemoryGB = 0; bool PreferNVLink = false; bool PreferSameNUMA = false; std::string Architecture; };
struct TopologyHint { std::string PreferredNode; std::vector<std::string> PreferredGPUs; int Score = 0; double BandwidthEstimate = 0; std::vector<std::string> Reasons; };

// kube-scheduler extender wire types (k8s.io/kube-scheduler/extender/v1: ExtenderFilterResult, HostPriority,
// MaxExtenderPriority = 10).  The reference only CONFIGURES an extender (verbs filter / prioritize,
// deploy/helm/kgwe/templates/scheduler-configmap.yaml:61-79) and ships no handler; these are the payloads
// such handlers fill from the engine ("next" row 8f.2); the HTTP/JSON transport around them is kgwe_extender.hpp.
struct ExtenderFilterResult { std::vector<std::string> NodeNames; std::map<std::string, std::string> FailedNodes; std::string Error; };
struct HostPriority { std::string Host; int64_t Score = 0; };
constexpr int64_t MaxExtenderPriority = 10;

// Go-style (value, error) result: ok() == (err == nil).
template <typename T>
struct Result { T value; std::string error; bool ok() const { return error.empty(); } };

// ---- packer: NodeTopology -> kgwe_node_image ----
struct PackedNode { kgwe_node_image image; std::vector<std::string> uuids; };
uint8_t ArchCode(const std::string& arch);                 // "" -> ANY, known names -> KGWE_ARCH_*, others interned 7..254
uint8_t MIGProfileCode(const std::string& name);           // 1..10 by the CRD whitelist, else KGWE_MIG_UNKNOWN
uint8_t TopologyPreferenceCode(const std::string& pref);
// Returns an error string ("" on success): more than 8 GPUs, more than 255 free instances of one profile, ...
std::string PackNode(const NodeTopology& node, PackedNode* out);

class TopologyAwareScheduler {
public:
    // NewTopologyAwareScheduler (scheduler.go:97-111); discovery is replaced by UpsertNode/RemoveNode.
    static Result<std::unique_ptr<TopologyAwareScheduler>> New(const SchedulerConfig& cfg, uint32_t maxNodes, int device = 0);
    ~TopologyAwareScheduler();

    // informer events (replaces RefreshTopology's full rebuild, discovery.go:290-319, 589-600)
    std::string UpsertNode(const NodeTopology& node);
    std::string RemoveNode(const std::string& nodeName);

    // Schedule (scheduler.go:114-179).  hint = the optimizer's PlacementHint, may be null.
    Result<SchedulingDecision> Schedule(const GPUWorkload& workload, const PlacementHint* hint = nullptr);
    // The same for a burst, one engine call (commits in order).
    std::vector<Result<SchedulingDecision>> ScheduleBurst(const std::vector<GPUWorkload>& workloads);
    // scoreNodes + descending sort (scheduler.go:137-146), best `topk` nodes.
    Result<std::vector<NodeScore>> ScoreNodes(const GPUWorkload& workload, const PlacementHint* hint, uint32_t topk);
    std::string ReleaseAllocation(const std::string& workloadUID);      // scheduler.go:710-727
    // DiscoveryService.GetTopologyHint (discovery.go:222-247) and the exporter's per-node gauge (prometheus_exporter.go:517-539)
    Result<TopologyHint> GetTopologyHint(const HintRequirements& req);
    // Extender verbs for one pod over the candidate node names kube-scheduler sends (nodeCacheCapable: false).
    // filter: keeps the nodes scoreNodes would keep (isNodeEligible and TotalScore > 0, scheduler.go:189-200);
    // prioritize: TotalScore clamped to 0..100 and scaled to 0..MaxExtenderPriority.
    ExtenderFilterResult Filter(const GPUWorkload& workload, const std::vector<std::string>& nodeNames);
    std::vector<HostPriority> Prioritize(const GPUWorkload& workload, const std::vector<std::string>& nodeNames);
    // The extender's bind verb for one pod: score it once more, and if `nodeName` still keeps it, record the GPUs the scorer
    // selects there in the ledger (kgwe_alloc_commit refuses GPUs another workload took in the meantime,
    // tryScheduleOnNode's re-check, scheduler.go:634-640).  Errors: the reference's strings where it has one.
    Result<SchedulingDecision> Bind(const GPUWorkload& workload, const std::string& nodeName);
    // Prometheus exposition text of the scheduler counters (prometheus_exporter.go:258-290 names) and of the engine's latency
    // histogram (kgwe_stats.latency_us_bucket: the reference's bucket edges, in microseconds).
    std::string MetricsText();
    Result<std::map<std::string, int>> TopologyScores();
    SchedulerMetrics GetMetrics() const { return metrics_; }            // scheduler.go:793
    std::deque<SchedulingEvent>& Events() { return events_; }           // scheduler.go:800 (a queue instead of a channel)

    bool HasTopology() const { return !node_names_.empty(); }
    uint32_t NodeIndex(const std::string& name) const;                  // KGWE_NO_NODE if unknown

private:
    TopologyAwareScheduler() = default;
    kgwe_pod_req PodReq(const GPUWorkload& w, bool commits);
    std::vector<uint8_t> HintBitmap(const PlacementHint* hint) const;
    NodeScore NodeScoreOf(const kgwe_placement& pl) const;
    SchedulingDecision DecisionOf(const GPUWorkload& w, const kgwe_placement& pl) const;
    void UpdateLatency(double ms);
    std::string ScoreAll(const GPUWorkload& w, std::vector<kgwe_node_score>* scores);

    SchedulerConfig cfg_;
    kgwe_engine* engine_ = nullptr;
    uint32_t max_nodes_ = 0;
    std::unordered_map<std::string, uint32_t> node_index_;
    std::vector<std::string> node_names_;                 // index -> name ("" = removed)
    std::vector<std::vector<std::string>> node_uuids_;    // index -> GPU UUIDs in node.GPUs order
    std::unordered_map<std::string, uint64_t> wl_id_;     // workload UID -> ledger key
    uint64_t next_wl_ = 1;
    SchedulerMetrics metrics_;
    std::deque<SchedulingEvent> events_;
};

}  // namespace kgwe_host
