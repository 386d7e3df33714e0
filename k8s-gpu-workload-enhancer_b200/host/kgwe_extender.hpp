// kgwe_extender.hpp — the kube-scheduler extender front the reference's Helm chart configures but never implements:
// verbs filter / prioritize / bind over HTTP + JSON (deploy/helm/kgwe/templates/scheduler-configmap.yaml:61-79:
// urlPrefix http://<release>-controller.<ns>.svc:8080, nodeCacheCapable false, managed resources nvidia.com/gpu and
// nvidia.com/mig-*), in front of the C++ mirror of TopologyAwareScheduler (kgwe_host.hpp) and through it the engine.
//
// Wire types are those of k8s.io/kube-scheduler/extender/v1 (a module the reference does not pin):
//   POST /filter      ExtenderArgs {pod, nodes | nodenames}  -> ExtenderFilterResult {nodes | nodenames, failedNodes, error}
//   POST /prioritize  ExtenderArgs                           -> HostPriorityList [{host, score 0..10}]
//   POST /bind        ExtenderBindingArgs {podName, podNamespace, podUID, node} -> ExtenderBindingResult {error}
//   GET  /metrics     Prometheus text: scheduler counters + the latency histogram of kgwe_stats
//   GET  /healthz
// A pod becomes a GPUWorkload like this (the reference has no conversion code; field meaning from the CRD,
// deploy/helm/kgwe/crds/gpuworkload-crd.yaml:56-122, resource names from the extender's managedResources):
//   GPUCount           sum over containers of resources.limits["nvidia.com/gpu"]
//   MIGRequirements    resources.limits["nvidia.com/mig-<profile>"] (first one found; Count = the quantity)
//   annotations        kgwe.io/topology-preference, kgwe.io/min-memory-gb, kgwe.io/gpu-architecture,
//                      kgwe.io/min-bandwidth-gbps
//   UID / Name / Namespace from metadata
// bind commits the GPUs the scorer selects on the named node (kgwe_alloc_commit, with its conflict check) for the pod seen
// last under that UID in filter / prioritize; the Binding object itself is the API server's business, not the engine's.
#pragma once
#include <atomic>
#include <map>
#include <string>
#include <vector>

#include "kgwe_host.hpp"

namespace kgwe_host {

// ---- a small JSON value (objects keep insertion order; look-ups ignore case like encoding/json does) ----
struct Json {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false; double num = 0; std::string str;
    std::vector<Json> items;                                   // Array
    std::vector<std::pair<std::string, Json>> fields;          // Object
    const Json* get(const std::string& key) const;             // case-insensitive; nullptr if absent or not an object
    static bool Parse(const std::string& text, Json* out, std::string* err);
    static std::string Quote(const std::string& s);
};

// Pod JSON (v1.Pod) -> GPUWorkload, by the rules above.  Returns "" or an error.
std::string WorkloadFromPod(const Json& pod, GPUWorkload* out);

class ExtenderServer {
public:
    explicit ExtenderServer(TopologyAwareScheduler* s) : sched_(s) {}
    // The three verbs on already-parsed bodies (also what the tests call directly): request JSON text -> response JSON text.
    std::string HandleFilter(const std::string& body);
    std::string HandlePrioritize(const std::string& body);
    std::string HandleBind(const std::string& body);
    std::string MetricsText();
    // Blocking accept loop on 0.0.0.0:port (port 0 = pick one).  on_listen, if given, is called with the bound port once the
    // socket listens.  One request per connection.  Returns when Stop() is called (from another thread, or by
    // POST /quitquitquit) or on a socket error.
    std::string Serve(int port, void (*on_listen)(int) = nullptr);
    void Stop() { stop_ = true; }

private:
    std::string Route(const std::string& method, const std::string& path, const std::string& body, int* status, std::string* ctype);
    bool NodeNamesOf(const Json& args, std::vector<std::string>* names, bool* as_nodes) const;
    TopologyAwareScheduler* sched_;
    std::map<std::string, GPUWorkload> seen_;     // pod UID -> the workload of its last filter / prioritize call (bind only carries names)
    std::atomic<bool> stop_{false};
    int listen_fd_ = -1;
};

}  // namespace kgwe_host
