// kgwe_cgo.go — the cgo shim a maintainer drops into the reference's src/scheduler package to put
// the B200 engine behind TopologyAwareScheduler (see INTEGRATION.md).
//
// NOT COMPILED IN THIS REPOSITORY: the build image has no Go toolchain (go / gccgo / tinygo are
// absent, here and on the GPU box).  It is kept as a thin 1:1 mapping over include/kgwe_engine.h
// so that it can be reviewed by reading; the same entry points, call for call, are exercised by
// the C++ mirror next to this file (kgwe_host.cpp, tests/host/) and by the ctypes binding.
// Last checked against include/kgwe_engine.h at KGWE_ABI_VERSION 4 (kgwe_node_upsert_remap,
// KGWE_ERR_CONFLICT from kgwe_alloc_commit, candidates_resolved and the latency histogram in
// kgwe_stats; version 4 only added the multi-GPU kgwe_schedule_queue_p2p, not bound here); init() below refuses to run
// against any other revision.
//
// Wiring inside the reference (src/scheduler/scheduler.go):
//   NewTopologyAwareScheduler  -> also newGPUEngine(config, maxNodes)
//   discovery informer events  -> engine.upsertNode(node) / engine.removeNode(name)
//   Schedule (:114-179)        -> engine.schedule(workload, hint) instead of scoreNodes + sort + tryScheduleOnNode + recordAllocation
//   ReleaseAllocation (:710)   -> engine.release(uid)
// Types GPUWorkload / SchedulingDecision / NodeScore and Schedule's signature stay as they are.
// The scheduler image must be built with CGO_ENABLED=1 (today: CGO_ENABLED=0 on scratch,
// docker/Dockerfile.scheduler:22,33) and carry libkgwe_engine.so plus the NVIDIA driver libraries.

//go:build cgo && kgwe_gpu

package scheduler

/*
#cgo CFLAGS:  -I${SRCDIR}/../../third_party/kgwe/include
#cgo LDFLAGS: -L${SRCDIR}/../../third_party/kgwe/lib -lkgwe_engine
#include <stdlib.h>
#include "kgwe_engine.h"
*/
import "C"

import (
	"fmt"
	"sort"
	"sync"
	"unsafe"

	"github.com/nvidia/kgwe/src/discovery"
)

const abiChecked = 4 // the KGWE_ABI_VERSION this file mirrors

func init() {
	if C.KGWE_ABI_VERSION != abiChecked || C.sizeof_kgwe_stats != 200 {
		panic(fmt.Sprintf("kgwe_cgo.go mirrors ABI %d of kgwe_engine.h, the header is at %d (%s)", abiChecked, int(C.KGWE_ABI_VERSION), C.GoString(C.kgwe_version())))
	}
}

var migNames = []string{"1g.5gb", "1g.10gb", "2g.10gb", "2g.20gb", "3g.20gb", "3g.40gb", "4g.20gb", "4g.40gb", "7g.40gb", "7g.80gb"}

// gpuEngine owns one kgwe_engine (one GPU, one node shard) plus the string <-> index tables.
// Strings never cross the C ABI.
type gpuEngine struct {
	mu        sync.Mutex
	h         *C.kgwe_engine
	nodeIdx   map[string]uint32 // node name -> dense index
	nodeName  []string          // index -> name ("" = removed)
	nodeUUIDs [][]string        // index -> GPU UUIDs in node.GPUs order
	wlID      map[string]uint64 // workload UID -> ledger key
	nextWL    uint64
	archExtra map[string]uint8 // architectures beyond the six known ones, interned 7..254
}

func newGPUEngine(cfg SchedulerConfig, maxNodes int) (*gpuEngine, error) {
	var c C.kgwe_config
	C.kgwe_config_default(&c)
	c.max_nodes = C.uint32_t(maxNodes)
	c.topology_weight = C.int32_t(cfg.TopologyWeight)
	c.resource_weight = C.int32_t(cfg.ResourceWeight)
	c.balance_weight = C.int32_t(cfg.BalanceWeight)
	e := &gpuEngine{nodeIdx: map[string]uint32{}, wlID: map[string]uint64{}, nextWL: 1, archExtra: map[string]uint8{}}
	if rc := C.kgwe_engine_create(&c, &e.h); rc != C.KGWE_OK {
		return nil, fmt.Errorf("kgwe_engine_create: %d (no CUDA device? the engine has no CPU fallback)", int(rc))
	}
	return e, nil
}

func (e *gpuEngine) close() { C.kgwe_engine_destroy(e.h) }

func (e *gpuEngine) err(rc C.int) error {
	return fmt.Errorf("kgwe: %d: %s", int(rc), C.GoString(C.kgwe_last_error(e.h)))
}

func (e *gpuEngine) archCode(a discovery.GPUArchitecture) uint8 {
	switch a {
	case "":
		return C.KGWE_ARCH_ANY
	case discovery.ArchitectureVolta:
		return C.KGWE_ARCH_VOLTA
	case discovery.ArchitectureTuring:
		return C.KGWE_ARCH_TURING
	case discovery.ArchitectureAmpere:
		return C.KGWE_ARCH_AMPERE
	case discovery.ArchitectureHopper:
		return C.KGWE_ARCH_HOPPER
	case discovery.ArchitectureAda:
		return C.KGWE_ARCH_ADA
	case discovery.ArchitectureBlackwell:
		return C.KGWE_ARCH_BLACKWELL
	}
	if c, ok := e.archExtra[string(a)]; ok {
		return c
	}
	if len(e.archExtra) >= 247 {
		return 255
	}
	c := uint8(7 + len(e.archExtra))
	e.archExtra[string(a)] = c
	return c
}

func migCode(name string) uint8 {
	for i, n := range migNames {
		if n == name {
			return uint8(i + 1)
		}
	}
	return C.KGWE_MIG_UNKNOWN
}

func prefCode(p TopologyPreference) uint8 {
	switch p {
	case TopologyPreferenceNVLinkOptimal:
		return C.KGWE_TOPO_NVLINK_OPTIMAL
	case TopologyPreferenceNVLinkRequired:
		return C.KGWE_TOPO_NVLINK_REQUIRED
	case TopologyPreferenceSameNUMA:
		return C.KGWE_TOPO_SAME_NUMA
	case TopologyPreferenceSamePCIeSwitch:
		return C.KGWE_TOPO_SAME_PCIE_SWITCH
	}
	return C.KGWE_TOPO_NONE // "None" and anything else: the default arm, scheduler.go:325-331
}

// rankCompress maps NUMANode values to small order-preserving ranks (only order and equality
// matter to scoreNUMATopology, scheduler.go:438-472).
func rankCompress(gpus []*discovery.GPUDevice) []int8 {
	vals := map[int]bool{}
	for _, g := range gpus {
		vals[g.Topology.NUMANode] = true
	}
	sorted := make([]int, 0, len(vals))
	for v := range vals {
		sorted = append(sorted, v)
	}
	sort.Ints(sorted)
	rank := map[int]int8{}
	for i, v := range sorted {
		rank[v] = int8(i)
	}
	out := make([]int8, len(gpus))
	for i, g := range gpus {
		out[i] = rank[g.Topology.NUMANode]
	}
	return out
}

// upsertNode packs one NodeTopology (discovery/types.go:351-366) into a kgwe_node_image.  Called
// from the informer handler instead of RefreshTopology's full rebuild (discovery.go:290-319).
func (e *gpuEngine) upsertNode(n *discovery.NodeTopology) error {
	e.mu.Lock()
	defer e.mu.Unlock()
	if len(n.GPUs) > C.KGWE_MAX_GPUS {
		return fmt.Errorf("node %s: more than %d GPUs", n.NodeName, C.KGWE_MAX_GPUS)
	}
	idx, ok := e.nodeIdx[n.NodeName]
	if !ok {
		idx = uint32(len(e.nodeName))
		e.nodeIdx[n.NodeName] = idx
		e.nodeName = append(e.nodeName, n.NodeName)
		e.nodeUUIDs = append(e.nodeUUIDs, nil)
	}
	var img C.kgwe_node_image
	gpuOf := map[string]int{}
	uuids := make([]string, len(n.GPUs))
	for i, g := range n.GPUs {
		gpuOf[g.UUID] = i
		uuids[i] = g.UUID
	}
	switchID := map[string]uint8{}
	numaRank := rankCompress(n.GPUs)
	img.n_gpus = C.uint8_t(len(n.GPUs))
	if n.NVSwitchInfo != nil && n.NVSwitchInfo.Present {
		img.node_flags |= C.KGWE_NODE_NVSWITCH
	}
	for i, g := range n.GPUs {
		if len(g.Topology.NVLinkConnections) > 0 {
			img.node_flags |= C.KGWE_NODE_ANY_NVLINK
		}
		for peer, l := range g.Topology.NVLinkConnections {
			j, ok := gpuOf[peer]
			if !ok || j == i {
				continue // links to GPUs of other nodes never reach the scorer
			}
			img.nvl_bw[i][j] = C.double(l.BandwidthGBps)
			img.nvl_present[i] |= 1 << uint(j)
			if l.IsActive {
				img.nvl_active[i] |= 1 << uint(j)
			}
		}
		img.pcie_bw[i] = C.double(g.Topology.PCIeInfo.BandwidthGBps)
		img.free_bytes[i] = C.uint64_t(g.Memory.FreeBytes)
		img.util[i] = C.double(g.Utilization.GPUPercent)
		sw := g.Topology.PCIeInfo.SwitchID
		if sw == "" {
			sw = "default" // scheduler.go:484-486
		}
		if _, ok := switchID[sw]; !ok {
			switchID[sw] = uint8(len(switchID)) // first-appearance order over node.GPUs
		}
		img.pcie_switch[i] = C.uint8_t(switchID[sw])
		img.numa[i] = C.int8_t(numaRank[i])
		img.arch[i] = C.uint8_t(e.archCode(g.Architecture))
		if g.Health.Status == discovery.HealthStatusHealthy {
			img.healthy_mask |= 1 << uint(i)
		}
		if g.MIGConfig != nil && g.MIGConfig.Enabled {
			img.mig_enabled_mask |= 1 << uint(i)
			for _, inst := range g.MIGConfig.CurrentInstances {
				if inst.InUse {
					continue
				}
				img.mig_has_free_mask |= 1 << uint(i)
				if p := migCode(inst.Profile.Name); p >= 1 && p <= C.KGWE_MIG_PROFILES {
					if img.mig_free[p-1] == 255 { // one byte per profile in the record: refuse, do not clamp
						return fmt.Errorf("node %s: more than 255 free MIG instances of profile %s", n.NodeName, inst.Profile.Name)
					}
					img.mig_free[p-1]++ // scheduler.go:606-622
				}
			}
		}
	}
	// The reference keys allocations by GPU UUID (scheduler.go:587-593), so a refresh that enumerates the node's GPUs in
	// another order keeps them; the engine's ledger holds index masks, so it is told where each old index went.
	var oldToNew [C.KGWE_MAX_GPUS]C.int8_t
	moved := false
	for g := range oldToNew {
		oldToNew[g] = -1
		if g < len(e.nodeUUIDs[idx]) {
			if q, ok := gpuOf[e.nodeUUIDs[idx][g]]; ok {
				oldToNew[g] = C.int8_t(q)
			}
			moved = moved || int(oldToNew[g]) != g
		}
	}
	var rc C.int
	if ok && moved {
		rc = C.kgwe_node_upsert_remap(e.h, C.uint32_t(idx), &img, &oldToNew[0])
	} else {
		rc = C.kgwe_node_upsert(e.h, C.uint32_t(idx), &img)
	}
	if rc != C.KGWE_OK {
		return e.err(rc)
	}
	e.nodeUUIDs[idx] = uuids
	return nil
}

func (e *gpuEngine) removeNode(name string) error {
	e.mu.Lock()
	defer e.mu.Unlock()
	idx, ok := e.nodeIdx[name]
	if !ok {
		return fmt.Errorf("node %s not found", name)
	}
	if rc := C.kgwe_node_remove(e.h, C.uint32_t(idx)); rc != C.KGWE_OK {
		return e.err(rc)
	}
	delete(e.nodeIdx, name) // the index stays retired; a re-added node gets a fresh one
	e.nodeName[idx] = ""
	return nil
}

func (e *gpuEngine) idOf(uid string) uint64 {
	id, ok := e.wlID[uid]
	if !ok {
		id = e.nextWL
		e.nextWL++
		e.wlID[uid] = id
	}
	return id
}

func (e *gpuEngine) podReq(w *GPUWorkload) C.kgwe_pod_req {
	var p C.kgwe_pod_req
	p.wl_id = C.uint64_t(e.idOf(w.UID))
	p.gpu_count = C.int32_t(w.Requirements.GPUCount)
	p.min_memory_gb = C.int32_t(w.Requirements.MinMemoryGB)
	p.min_bandwidth_gbps = 0 // reference behaviour; set from MinBandwidthGBps to enable extension E3
	p.topology_pref = C.uint8_t(prefCode(w.Requirements.TopologyPreference))
	p.arch = C.uint8_t(e.archCode(w.Requirements.GPUArchitecture))
	if m := w.Requirements.MIGRequirements; m != nil {
		p.mig_profile = C.uint8_t(migCode(m.Profile)) // 255 when the name is not in the table
		p.mig_count = C.int32_t(m.Count)
	}
	return p
}

// hintBitmap: bit n set when node n is in PlacementHint.PreferredNodes (scheduler.go:270-279).
func (e *gpuEngine) hintBitmap(preferred []string) []byte {
	bm := make([]byte, (len(e.nodeName)+7)/8)
	for _, n := range preferred {
		if i, ok := e.nodeIdx[n]; ok {
			bm[i/8] |= 1 << (i & 7)
		}
	}
	return bm
}

func orderList(order C.uint32_t) []int {
	n := int(order>>24) & 15
	out := make([]int, n)
	for p := 0; p < n; p++ {
		out[p] = int(order>>(3*uint(p))) & 7
	}
	return out
}

func (e *gpuEngine) nodeScoreOf(pl C.kgwe_placement) NodeScore {
	s := NodeScore{
		NodeName:      e.nodeName[pl.node_idx],
		TotalScore:    int(pl.total),
		TopologyScore: int(pl.topology),
		ResourceScore: int(pl.resource),
		BalanceScore:  int(pl.balance),
		Reasons:       make([]string, 0),
	}
	for _, g := range orderList(pl.order) {
		s.SelectedGPUs = append(s.SelectedGPUs, e.nodeUUIDs[pl.node_idx][g])
	}
	if pl.flags&C.KGWE_F_HINT != 0 { // scheduler.go:270-297
		s.Reasons = append(s.Reasons, "ML optimizer preferred this node")
	}
	if s.TopologyScore >= 80 {
		s.Reasons = append(s.Reasons, "Optimal NVLink connectivity available")
	}
	if s.ResourceScore >= 80 {
		s.Reasons = append(s.Reasons, "Sufficient resources with headroom")
	}
	if s.BalanceScore >= 80 {
		s.Reasons = append(s.Reasons, "Good load distribution")
	}
	return s
}

// scoreNodes replacement (scheduler.go:182-203 + the sort at :144-146): best `topk` nodes, ranked.
func (e *gpuEngine) scoreNodes(w *GPUWorkload, hint *PlacementHint, topk int) ([]NodeScore, error) {
	e.mu.Lock()
	defer e.mu.Unlock()
	p := e.podReq(w)
	out := make([]C.kgwe_placement, topk)
	var hintPtr *C.uint8_t
	if hint != nil && len(e.nodeName) > 0 {
		bm := e.hintBitmap(hint.PreferredNodes)
		hintPtr = (*C.uint8_t)(unsafe.Pointer(&bm[0]))
	}
	if rc := C.kgwe_score_batch(e.h, &p, 1, hintPtr, nil, C.uint32_t(topk), &out[0]); rc != C.KGWE_OK {
		return nil, e.err(rc)
	}
	scores := make([]NodeScore, 0, topk)
	for _, pl := range out {
		if pl.node_idx == C.KGWE_NO_NODE {
			break
		}
		scores = append(scores, e.nodeScoreOf(pl))
	}
	return scores, nil
}

// schedule: Schedule's body after the snapshot check and the optimizer call (scheduler.go:136-167).
func (e *gpuEngine) schedule(w *GPUWorkload, hint *PlacementHint) (*SchedulingDecision, NodeScore, error) {
	e.mu.Lock()
	defer e.mu.Unlock()
	p := e.podReq(w)
	var pl C.kgwe_placement
	var hintPtr *C.uint8_t
	if hint != nil && len(e.nodeName) > 0 {
		bm := e.hintBitmap(hint.PreferredNodes)
		hintPtr = (*C.uint8_t)(unsafe.Pointer(&bm[0]))
	}
	if rc := C.kgwe_schedule_queue(e.h, &p, 1, hintPtr, &pl); rc != C.KGWE_OK {
		return nil, NodeScore{}, e.err(rc)
	}
	if pl.node_idx == C.KGWE_NO_NODE {
		return nil, NodeScore{}, fmt.Errorf("no eligible nodes found for workload %s", w.Name) // scheduler.go:140
	}
	ns := e.nodeScoreOf(pl)
	return &SchedulingDecision{
		WorkloadUID:        w.UID,
		SelectedNode:       ns.NodeName,
		SelectedGPUs:       ns.SelectedGPUs,
		Score:              ns.TotalScore,
		Reasons:            ns.Reasons,
		EstimatedBandwidth: float64(pl.est_bandwidth_gbps),
	}, ns, nil
}

// scheduleBurst: the same for a queue of workloads, one engine call, commits in order.
func (e *gpuEngine) scheduleBurst(ws []*GPUWorkload) ([]*SchedulingDecision, error) {
	e.mu.Lock()
	defer e.mu.Unlock()
	if len(ws) == 0 {
		return nil, nil
	}
	pods := make([]C.kgwe_pod_req, len(ws))
	for i, w := range ws {
		pods[i] = e.podReq(w)
	}
	out := make([]C.kgwe_placement, len(ws))
	if rc := C.kgwe_schedule_queue(e.h, &pods[0], C.uint32_t(len(ws)), nil, &out[0]); rc != C.KGWE_OK {
		return nil, e.err(rc)
	}
	ds := make([]*SchedulingDecision, len(ws))
	for i, pl := range out {
		if pl.node_idx == C.KGWE_NO_NODE {
			continue // the caller reports "no eligible nodes found for workload %s"
		}
		ns := e.nodeScoreOf(pl)
		ds[i] = &SchedulingDecision{WorkloadUID: ws[i].UID, SelectedNode: ns.NodeName, SelectedGPUs: ns.SelectedGPUs,
			Score: ns.TotalScore, Reasons: ns.Reasons, EstimatedBandwidth: float64(pl.est_bandwidth_gbps)}
	}
	return ds, nil
}

// release: ReleaseAllocation (scheduler.go:710-727).
func (e *gpuEngine) release(uid string) error {
	e.mu.Lock()
	defer e.mu.Unlock()
	id, ok := e.wlID[uid]
	if !ok || C.kgwe_alloc_release(e.h, C.uint64_t(id)) != C.KGWE_OK {
		return fmt.Errorf("allocation not found for workload %s", uid) // scheduler.go:715
	}
	return nil
}

// bind: the extender's bind verb for one pod (deploy/helm/kgwe/templates/scheduler-configmap.yaml:64): score the pod
// once more, and if nodeName still keeps it, record the GPUs the scorer selects there.  kgwe_alloc_commit refuses GPUs
// another workload took in the meantime (KGWE_ERR_CONFLICT), which is tryScheduleOnNode's re-check (scheduler.go:634-640).
func (e *gpuEngine) bind(w *GPUWorkload, nodeName string) (*SchedulingDecision, error) {
	e.mu.Lock()
	defer e.mu.Unlock()
	idx, ok := e.nodeIdx[nodeName]
	if !ok {
		return nil, fmt.Errorf("node %s not found", nodeName)
	}
	n := len(e.nodeName)
	req := e.podReq(w)
	scores := make([]C.kgwe_node_score, n)
	if rc := C.kgwe_score_batch(e.h, &req, 1, nil, &scores[0], C.uint32_t(n), 0, nil); rc != C.KGWE_OK {
		return nil, e.err(rc)
	}
	s := scores[idx]
	if s.flags&C.KGWE_F_KEPT == 0 {
		return nil, fmt.Errorf("failed to schedule on any eligible node") // scheduler.go:178
	}
	switch rc := C.kgwe_alloc_commit(e.h, C.uint64_t(e.idOf(w.UID)), C.uint32_t(idx), s.gpu_mask); rc {
	case C.KGWE_OK:
	case C.KGWE_ERR_CONFLICT:
		return nil, fmt.Errorf("GPU already allocated on node %s", nodeName)
	default:
		return nil, e.err(rc)
	}
	pl := C.kgwe_placement{node_idx: C.uint32_t(idx), total: s.total, topology: s.topology, resource: s.resource, balance: s.balance, gpu_mask: s.gpu_mask, flags: s.flags, order: s.order}
	ns := e.nodeScoreOf(pl)
	return &SchedulingDecision{WorkloadUID: w.UID, SelectedNode: ns.NodeName, SelectedGPUs: ns.SelectedGPUs, Score: ns.TotalScore, Reasons: ns.Reasons}, nil
}

// schedulerMetrics: GetMetrics (scheduler.go:793) plus the engine's own counters.  The latency histogram has the bucket
// edges of src/monitoring/prometheus_exporter.go:262 read as MICROSECONDS (10 us ... 5 ms, +Inf).
func (e *gpuEngine) schedulerMetrics() (SchedulerMetrics, [C.KGWE_LATENCY_BUCKETS]uint64, error) {
	var st C.kgwe_stats
	var buckets [C.KGWE_LATENCY_BUCKETS]uint64
	if rc := C.kgwe_get_stats(e.h, &st); rc != C.KGWE_OK {
		return SchedulerMetrics{}, buckets, e.err(rc)
	}
	for i := range buckets {
		buckets[i] = uint64(st.latency_us_bucket[i])
	}
	m := SchedulerMetrics{TotalScheduled: int64(st.scheduled), TotalFailed: int64(st.failed), TopologyOptimalPlacements: int64(st.topology_optimal)}
	if st.latency_count > 0 {
		m.AverageLatencyMs = float64(st.latency_sum_us) / float64(st.latency_count) / 1000.0
	}
	return m, buckets, nil
}
