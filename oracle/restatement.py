"""oracle/restatement.py — TEST INFRASTRUCTURE (nothing under k8s-gpu-workload-enhancer_b200/ imports it).

A third restatement of the reference's Filter / Prioritize / Schedule path (and, below, of DiscoveryService.GetTopologyHint,
src/discovery/discovery.go:222-247, 378-541, and of the exporter's per-node gauge, src/monitoring/prometheus_exporter.go:
517-539), written from /root/reference/src/scheduler/scheduler.go:114-707 on its own - not from oracle/kgwe_oracle.cpp and not from the
kernels - in plain Python with the reference's own data shapes: GPU objects with UUID strings, per-call adjacency
dicts, a dict ledger scanned per node.  tests/test_restatements_agree.py runs it against both builds of the C++
oracle on small random clusters; three texts that agree make a transcription slip in any one of them visible.
(The reference has no tests and cannot run here - no Go toolchain - so parity stays "unpinned by reference tests";
this narrows what that can hide.)

Only reference semantics: no extension (E1 exhaustive, E3 minimum bandwidth) is restated here.

Canonical determinisation, the same legal outcomes of Go's map iteration the other two restatements pick
(SURVEY.md 8c): nodes in index order, final pick = highest TotalScore then lowest node index, NUMA groups in
ascending NUMANode, PCIe groups in ascending interned switch id.
"""
from dataclasses import dataclass, field

NVLINK_PREFS = ("NVLinkRequired", "NVLinkOptimal")


@dataclass
class GPU:
    uuid: str
    architecture: int
    free_bytes: int
    gpu_percent: float
    healthy: bool
    numa_node: int
    switch_id: int                     # interned PCIe switch id ("" -> "default" happens in the packer)
    pcie_bandwidth: float
    nvlink: dict = field(default_factory=dict)      # peer uuid -> (BandwidthGBps, IsActive), in ascending peer index (the canonical map order)
    mig_enabled: bool = False
    mig_has_free: bool = False                      # some MIG instance of the GPU is not in use


@dataclass
class Node:
    name: str
    index: int
    gpus: list
    mig_free: list                     # free instances per profile over the node's MIG-enabled GPUs (getMIGAvailability's count)
    nvswitch_present: bool = False


@dataclass
class Workload:
    uid: int
    gpu_count: int
    min_memory_gb: int
    preference: str
    architecture: int                  # 0 = ""
    mig_profile: int                   # 0 = no MIGRequirements
    mig_count: int


@dataclass
class NodeScore:
    node: Node
    total: int = 0
    topology: int = 0
    resource: int = 0
    balance: int = 0
    selected: list = field(default_factory=list)
    hinted: bool = False


class Scheduler:
    """TopologyAwareScheduler with DefaultSchedulerConfig weights (types.go:379-392)."""

    def __init__(self, nodes, topology_weight=40, resource_weight=35, balance_weight=25):
        self.nodes = nodes                           # ClusterTopology.Nodes
        self.allocations = {}                        # workload UID -> (node name, [GPU UUIDs])
        self.wt, self.wr, self.wb = topology_weight, resource_weight, balance_weight

    # scheduler.go:581-603
    def get_available_gpus(self, node):
        allocated = set()
        for node_name, uuids in self.allocations.values():
            if node_name == node.name:
                allocated.update(uuids)
        return [g for g in node.gpus if g.uuid not in allocated and g.healthy]

    # scheduler.go:606-622 (the packer has already counted the free instances per profile)
    def get_mig_availability(self, node, profile):
        return node.mig_free[profile - 1] if 1 <= profile <= len(node.mig_free) else 0

    # scheduler.go:206-241
    def is_node_eligible(self, node, w):
        available = self.get_available_gpus(node)
        if len(available) < w.gpu_count:
            return False
        if w.architecture != 0:
            if not any(g.architecture == w.architecture for g in available):
                return False
        if w.mig_profile != 0:
            if self.get_mig_availability(node, w.mig_profile) < w.mig_count:
                return False
        return True

    # scheduler.go:376-435
    def find_best_nvlink_group(self, graph, available, size):
        if size <= 1 and available:
            return [available[0].uuid], 0.0
        best_group, best_bandwidth = [], 0.0
        for start in available:
            group = [start.uuid]
            total = 0.0
            while len(group) < size:
                best_candidate, best_candidate_bw = "", 0.0
                for cand in available:
                    if cand.uuid in group:
                        continue
                    agg, fully = 0.0, True
                    for member in group:
                        if member in graph[cand.uuid]:
                            agg += graph[cand.uuid][member]
                        else:
                            fully = False
                    if fully and agg > best_candidate_bw:
                        best_candidate, best_candidate_bw = cand.uuid, agg
                if best_candidate == "":
                    break
                group.append(best_candidate)
                total += best_candidate_bw
            if len(group) >= size and total > best_bandwidth:
                best_group, best_bandwidth = group[:size], total
        return best_group, best_bandwidth

    # scheduler.go:336-373
    def score_nvlink_topology(self, available, required):
        if len(available) < required:
            return 0, []
        graph = {}
        for g in available:
            graph[g.uuid] = {peer: bw for peer, (bw, active) in g.nvlink.items() if active}
        best_group, total = self.find_best_nvlink_group(graph, available, required)
        if len(best_group) < required:
            return 30, [g.uuid for g in available[:required]]
        max_bandwidth = 900.0 * float(required * (required - 1) // 2)
        ratio = total / max_bandwidth
        return int(50 + 50 * ratio), best_group

    # scheduler.go:438-472 / 475-513: the same shape, keyed by NUMANode / SwitchID
    @staticmethod
    def score_grouped(available, required, key, same_score, spread_score):
        groups = {}
        for g in available:
            groups.setdefault(key(g), []).append(g)
        for k in sorted(groups):                               # canonical order of the map iteration
            if len(groups[k]) >= required:
                return same_score, [g.uuid for g in groups[k][:required]]
        selected = []
        for k in sorted(groups):
            for g in groups[k]:
                selected.append(g.uuid)
                if len(selected) >= required:
                    return spread_score, selected
        return 0, []

    # scheduler.go:303-333
    def calculate_topology_score(self, available, w):
        required = w.gpu_count
        if required <= 1:
            if available:
                return 100, [available[0].uuid]
            return 0, []
        if w.preference in NVLINK_PREFS:
            return self.score_nvlink_topology(available, required)
        if w.preference == "SameNUMA":
            return self.score_grouped(available, required, lambda g: g.numa_node, 90, 50)
        if w.preference == "SamePCIeSwitch":
            return self.score_grouped(available, required, lambda g: g.switch_id, 80, 40)
        return 50, [g.uuid for g in available[:required]]

    # scheduler.go:516-553
    @staticmethod
    def calculate_resource_score(available, w):
        if not available:
            return 0
        score = 50
        total_free = sum(g.free_bytes for g in available) & 0xFFFFFFFFFFFFFFFF
        required_memory = (w.min_memory_gb * 1024 * 1024 * 1024 * w.gpu_count) & 0xFFFFFFFFFFFFFFFF
        if total_free >= (required_memory * 2) & 0xFFFFFFFFFFFFFFFF:
            score += 25
        elif total_free >= required_memory:
            score += 10
        avg = 0.0
        for g in available:
            avg += g.gpu_percent
        avg /= float(len(available))
        if avg < 30:
            score += 25
        elif avg < 60:
            score += 10
        return min(100, score)

    # scheduler.go:556-578
    def calculate_balance_score(self, node):
        node_allocations = sum(1 for node_name, _ in self.allocations.values() if node_name == node.name)
        total_gpus = len(node.gpus)
        if total_gpus == 0:
            return 0
        usage = float(node_allocations) / float(total_gpus)
        return max(0, int(100 * (1 - usage)))

    # scheduler.go:244-300
    def score_node(self, node, w, preferred):
        s = NodeScore(node)
        available = self.get_available_gpus(node)
        s.topology, s.selected = self.calculate_topology_score(available, w)
        s.resource = self.calculate_resource_score(available, w)
        s.balance = self.calculate_balance_score(node)
        weighted = s.topology * self.wt + s.resource * self.wr + s.balance * self.wb
        s.total = int(weighted / 100) if weighted < 0 else weighted // 100          # Go's integer division truncates toward zero
        if preferred is not None and node.name in preferred:
            s.total += 10
            s.hinted = True
        return s

    # scheduler.go:182-203: every node's score, None where the node is not eligible; `kept` = TotalScore > 0
    def score_nodes(self, w, preferred=None):
        out = []
        for node in self.nodes:
            out.append(self.score_node(node, w, preferred) if self.is_node_eligible(node, w) else None)
        return out

    # scheduler.go:656-692
    def estimate_bandwidth(self, s):
        if len(s.selected) <= 1:
            return 0.0
        total, pairs = 0.0, 0
        for i, u1 in enumerate(s.selected):
            for j, u2 in enumerate(s.selected):
                if i >= j:
                    continue
                for g in s.node.gpus:
                    if g.uuid == u1:
                        total += g.nvlink[u2][0] if u2 in g.nvlink else g.pcie_bandwidth
                        pairs += 1
        return 0.0 if pairs == 0 else total / float(pairs)

    # scheduler.go:114-179 without the optimizer call: score, sort (canonical), first fit, record
    def schedule(self, w, preferred=None):
        scores = [s for s in self.score_nodes(w, preferred) if s is not None and s.total > 0]
        if not scores:
            return None, 0.0
        scores.sort(key=lambda s: (-s.total, s.node.index))
        best = scores[0]                                   # tryScheduleOnNode cannot fail single-threaded: the GPUs were just seen free
        bandwidth = self.estimate_bandwidth(best)
        self.allocations[w.uid] = (best.node.name, list(best.selected))      # recordAllocation: same UID overwrites (:699)
        return best, bandwidth


# ---- ABI arrays (numpy records of k8s-gpu-workload-enhancer_b200/abi.py) -> the objects above ----------------------
PREF_NAMES = {0: "None", 1: "NVLinkOptimal", 2: "NVLinkRequired", 3: "SameNUMA", 4: "SamePCIeSwitch"}


def nodes_from_images(images):
    nodes = []
    for n, im in enumerate(images):
        gpus = []
        ng = int(im["n_gpus"])
        for g in range(ng):
            links = {}
            for p in range(ng):
                if int(im["nvl_present"][g]) >> p & 1:
                    links[f"n{n}g{p}"] = (float(im["nvl_bw"][g][p]), bool(int(im["nvl_active"][g]) >> p & 1))
            gpus.append(GPU(f"n{n}g{g}", int(im["arch"][g]), int(im["free_bytes"][g]), float(im["util"][g]),
                            bool(int(im["healthy_mask"]) >> g & 1), int(im["numa"][g]), int(im["pcie_switch"][g]),
                            float(im["pcie_bw"][g]), links, bool(int(im["mig_enabled_mask"]) >> g & 1), bool(int(im["mig_has_free_mask"]) >> g & 1)))
        nodes.append(Node(f"node{n}", n, gpus, [int(x) for x in im["mig_free"]], bool(int(im["node_flags"]) & 1)))
    return nodes


def workload_from_pod(pod):
    return Workload(int(pod["wl_id"]), int(pod["gpu_count"]), int(pod["min_memory_gb"]), PREF_NAMES.get(int(pod["topology_pref"]), "None"),
                    int(pod["arch"]), int(pod["mig_profile"]), int(pod["mig_count"]))


def seed_ledger(sched, ledger):
    for a in ledger:
        n, mask = int(a["node_idx"]), int(a["gpu_mask"])
        sched.allocations[int(a["wl_id"])] = (f"node{n}", [f"n{n}g{g}" for g in range(8) if mask >> g & 1])


def gpu_indices(selected):
    return [int(u.split("g")[1]) for u in selected]


# ---- DiscoveryService.GetTopologyHint (src/discovery/discovery.go:222-247, 378-541) --------------------------------
@dataclass
class HintRequirements:
    count: int
    min_memory_gb: int
    prefer_nvlink: bool
    prefer_same_numa: bool
    architecture: int                  # 0 = ""


class Discovery:
    def __init__(self, nodes):
        self.nodes = nodes

    # discovery.go:436-460
    @staticmethod
    def get_available_gpus(node):
        available = []
        for g in node.gpus:
            if g.healthy:
                if g.mig_enabled:
                    if g.mig_has_free:
                        available.append(g)
                elif g.gpu_percent < 90:
                    available.append(g)
        return available

    # discovery.go:462-488 (the peers of a GPU in ascending index: the canonical order of the map iteration at :468)
    @staticmethod
    def find_nvlink_groups(gpus, size):
        groups = []
        for g in gpus:
            group = [g.uuid]
            for peer_uuid in g.nvlink:
                if len(group) >= size:
                    break
                for peer in gpus:
                    if peer.uuid == peer_uuid:
                        group.append(peer_uuid)
                        break
            if len(group) >= size:
                groups.append(group[:size])
        return groups

    # discovery.go:490-504
    @staticmethod
    def are_on_same_numa(node, uuids):
        numa = -2
        for u in uuids:
            for g in node.gpus:
                if g.uuid == u:
                    if numa == -2:
                        numa = g.numa_node
                    elif numa != g.numa_node:
                        return False
        return True

    # discovery.go:506-541
    @staticmethod
    def estimate_bandwidth(node, uuids):
        if len(uuids) < 2:
            return 0.0
        total, count = 0.0, 0
        for i, u1 in enumerate(uuids):
            for j, u2 in enumerate(uuids):
                if i >= j:
                    continue
                for g in node.gpus:
                    if g.uuid == u1:
                        total += g.nvlink[u2][0] if u2 in g.nvlink else g.pcie_bandwidth
                        count += 1
                        break
        return 0.0 if count == 0 else total / float(count)

    # discovery.go:378-434
    def score_node_for_requirements(self, node, req):
        available = self.get_available_gpus(node)
        if len(available) < req.count:
            return 0, []
        filtered = [g for g in available if int(g.free_bytes // (1024 * 1024 * 1024)) >= req.min_memory_gb]
        if len(filtered) < req.count:
            return 0, []
        score, selected = 50, []
        if req.count > 1 and req.prefer_nvlink:
            groups = self.find_nvlink_groups(filtered, req.count)
            if groups:
                score += 30
                selected = groups[0]
        if not selected:
            selected = [g.uuid for g in filtered[:max(req.count, 0)]]
        if req.prefer_same_numa and self.are_on_same_numa(node, selected):
            score += 10
        if req.architecture != 0:
            for u in selected:
                for g in node.gpus:
                    if g.uuid == u and g.architecture == req.architecture:
                        score += 5
        return score, selected

    # discovery.go:222-247: (node index, score, GPU uuids, bandwidth estimate) or None ("no suitable placement found")
    def get_topology_hint(self, req):
        best = None
        best_score = 0
        for node in self.nodes:
            score, gpus = self.score_node_for_requirements(node, req)
            if score > best_score:
                best_score = score
                best = (node.index, score, gpus, self.estimate_bandwidth(node, gpus))
        return best


# PrometheusExporter.calculateTopologyScore (src/monitoring/prometheus_exporter.go:517-539)
def exporter_topology_score(node):
    if not node.gpus:
        return 0
    score = 50
    if node.nvswitch_present:
        score += 30
    if sum(len(g.nvlink) for g in node.gpus) > 0:
        score += 20
    return min(100, score)
