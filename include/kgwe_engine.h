/*
 * kgwe_engine.h — C ABI of the B200 scoring engine for KGWE's scheduler hot path.
 *
 * This is the drop-in boundary for ONE path of asklokesh/k8s-gpu-workload-enhancer:
 * the per-pod x per-node x per-GPU-subset Filter/Prioritize scoring done today by
 * (*TopologyAwareScheduler).scoreNodes and the argmax/commit around it in Schedule
 * (reference: src/scheduler/scheduler.go:114-179, 182-707).  A cgo shim in the
 * reference's src/scheduler package binds exactly these entry points (see
 * INTEGRATION.md); tests and bench.py bind the same symbols through ctypes.
 *
 * Conventions
 *  - Every function returns KGWE_OK (0) or a negative kgwe_status; nothing throws or
 *    aborts across this boundary.  kgwe_last_error() gives the message of the last
 *    failing call made BY THE CALLING THREAD (kept per thread, so another thread's
 *    failure cannot replace the text while it is being read).
 *  - All structs are plain old data with fixed-width fields.  Buffers are owned by
 *    the caller; the engine keeps no caller pointer after a call returns (cgo rule).
 *  - The engine owns its device memory and pinned staging.  One mutex per engine
 *    (mirrors `s.mu`, scheduler.go:17) makes every entry point callable from any
 *    thread; each call selects the engine's CUDA device first.
 *  - The cluster image is double-buffered on the device: node upserts and ledger calls
 *    only touch a host mirror; the next call that reads the image copies the pending
 *    records into the image that is NOT being read, on a copy stream, and switches over
 *    with event dependencies - no entry point synchronises the host with a stream to
 *    apply updates.  Calls given different cuda_stream arguments are ordered one after
 *    the other on the device the same way (the engine's scratch buffers are shared).
 *  - There is no CPU fallback: if CUDA is unavailable, kgwe_engine_create fails.
 *  - Canonical order: nodes are visited in ascending node index and
 *    ties on TotalScore resolve to the LOWEST node index; NUMA groups in ascending
 *    numa[] value; PCIe groups in ascending pcie_switch[] id.  Each is one of the
 *    outcomes Go's map iteration / unstable sort can produce.
 */
#ifndef KGWE_ENGINE_H
#define KGWE_ENGINE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KGWE_ABI_VERSION   4     /* bumped whenever an entry point or a struct of this header changes; kgwe_version() ends in "abi <n>" */
#define KGWE_MAX_GPUS      8     /* GPUs per node record (all BASELINE configs use 8) */
#define KGWE_MIG_PROFILES  10    /* CRD whitelist, src/sharing/mig_controller.go:273-284 */
#define KGWE_MAX_TOPK      16
#define KGWE_NO_NODE       0xFFFFFFFFu
#define KGWE_MAX_BW_GBPS   100000.0   /* upsert rejects bandwidths outside [0, this] or non-finite */
#define KGWE_LATENCY_BUCKETS 10  /* upper edges in microseconds: 10 25 50 100 250 500 1000 2500 5000 +Inf */

typedef enum kgwe_status {
    KGWE_OK                =  0,
    KGWE_ERR_INVALID_ARG   = -1,
    KGWE_ERR_CUDA          = -2,
    KGWE_ERR_NO_DEVICE     = -3,
    KGWE_ERR_OUT_OF_RANGE  = -4,
    KGWE_ERR_NOT_FOUND     = -5,   /* "allocation not found for workload", scheduler.go:715 */
    KGWE_ERR_CONFLICT      = -6,   /* commit names a GPU that is already allocated, scheduler.go:634-640 */
    KGWE_ERR_NO_MEMORY     = -7,
    KGWE_ERR_TOO_LARGE     = -8
} kgwe_status;

/* TopologyPreference, src/scheduler/types.go:60-77.  Any other string maps to NONE
 * (the `default:` arm, scheduler.go:325-331). */
enum {
    KGWE_TOPO_NONE = 0,
    KGWE_TOPO_NVLINK_OPTIMAL = 1,
    KGWE_TOPO_NVLINK_REQUIRED = 2,   /* treated exactly like OPTIMAL, scheduler.go:319-320 */
    KGWE_TOPO_SAME_NUMA = 3,
    KGWE_TOPO_SAME_PCIE_SWITCH = 4
};

/* GPUArchitecture, src/discovery/types.go:50-59.  0 = "" (no requirement / unknown). */
enum {
    KGWE_ARCH_ANY = 0, KGWE_ARCH_VOLTA = 1, KGWE_ARCH_TURING = 2, KGWE_ARCH_AMPERE = 3,
    KGWE_ARCH_HOPPER = 4, KGWE_ARCH_ADA = 5, KGWE_ARCH_BLACKWELL = 6
    /* 7..254: host-interned other strings; 255 = a requested string no GPU carries */
};

/* MIG profile ids: index+1 into the CRD whitelist in this order
 * (deploy/helm/kgwe/crds/gpuworkload-crd.yaml:94-104):
 * 1g.5gb 1g.10gb 2g.10gb 2g.20gb 3g.20gb 3g.40gb 4g.20gb 4g.40gb 7g.40gb 7g.80gb */
#define KGWE_MIG_NONE     0      /* MIGRequirements == nil */
#define KGWE_MIG_UNKNOWN  255    /* a profile name outside the table: availability is 0 */

/* Engine configuration.  Mirrors the fields of SchedulerConfig that the path reads
 * (src/scheduler/types.go:346-392; weights used at scheduler.go:282-284). */
typedef struct kgwe_config {
    uint32_t struct_size;        /* sizeof(kgwe_config), for forward compatibility */
    int32_t  device;             /* CUDA device ordinal */
    uint32_t max_nodes;          /* capacity of the cluster image (records) */
    uint32_t node_base;          /* global index of local record 0 (node sharding; 0 on one GPU) */
    int32_t  topology_weight;    /* 40  (types.go:385) */
    int32_t  resource_weight;    /* 35  (types.go:386) */
    int32_t  balance_weight;     /* 25  (types.go:387) */
    int32_t  hint_bonus;         /* 10  (scheduler.go:274) */
    double   nvlink_norm_gbps;   /* 900.0 (scheduler.go:368) */
    uint32_t flags;              /* KGWE_CFG_* */
    uint32_t reserved;
} kgwe_config;

/* kgwe_score_batch_device cannot look at the pods before launching: set this when device-resident
 * batches may contain KGWE_POD_EXHAUSTIVE pods, so the scoring kernel is launched with its per-node
 * subset tables (one CTA per SM instead of two).  Without it such pods are still scored correctly,
 * on a slower path.  kgwe_score_batch (host pods) picks the right kernel by itself. */
#define KGWE_CFG_SUBSET_TABLES 0x1u
/* The greedy group of a node depends only on the node's state and on GPUCount, not on the pod:
 * with this flag the scoring kernel finds it once per node tile for each count 2..8 and every
 * (pod,node) evaluation looks it up.  Results are identical; throughput figures obtained with it
 * are reported separately from the brute-force rate (the group search is shared between pods). */
#define KGWE_CFG_MEMOISE_GROUPS 0x2u
/* By default every available start GPU of a node grows its own group for every (pod,node) evaluation, as
 * findBestNVLinkGroup does (scheduler.go:389): A candidate groups scored per evaluation.  On a node whose
 * active links all carry the same bandwidth every group that reaches the requested size has the same total and
 * the reference keeps the first such start, so the walk from the lowest available start already is the
 * answer whenever it completes.  With this flag the scoring kernel stops there (and searches all starts only
 * when that walk dies).  Results are identical; kgwe_stats.candidates then counts the one group actually
 * grown, kgwe_stats.candidates_resolved still counts A. */
#define KGWE_CFG_PRUNE_TIED_STARTS 0x4u

/* One node of the topology snapshot: exactly the fields of NodeTopology / GPUDevice the
 * scorer reads (src/discovery/types.go:11-47,116-164,336-366).
 * 768 bytes.  GPU order = order of node.GPUs.  Entries at index >= n_gpus are ignored. */
typedef struct kgwe_node_image {
    double   nvl_bw[KGWE_MAX_GPUS][KGWE_MAX_GPUS]; /* [i][j] = GPU i's NVLinkConnections[uuid_j].BandwidthGBps; 0 if no entry */
    double   pcie_bw[KGWE_MAX_GPUS];               /* Topology.PCIeInfo.BandwidthGBps */
    uint64_t free_bytes[KGWE_MAX_GPUS];            /* Memory.FreeBytes */
    double   util[KGWE_MAX_GPUS];                  /* Utilization.GPUPercent */
    uint8_t  nvl_present[KGWE_MAX_GPUS];           /* bit j: GPU i has a NVLinkConnections entry for GPU j */
    uint8_t  nvl_active[KGWE_MAX_GPUS];            /* bit j: that entry has IsActive (subset of nvl_present) */
    int8_t   numa[KGWE_MAX_GPUS];                  /* Topology.NUMANode, order-preserving (packer rank-compresses) */
    uint8_t  pcie_switch[KGWE_MAX_GPUS];           /* PCIeInfo.SwitchID interned in first-appearance order, ""->"default" */
    uint8_t  arch[KGWE_MAX_GPUS];                  /* KGWE_ARCH_* */
    uint8_t  mig_free[KGWE_MIG_PROFILES];          /* per profile: # of !InUse instances on MIG-enabled GPUs (scheduler.go:606-622) */
    uint8_t  n_gpus;                               /* len(node.GPUs), 0..8 */
    uint8_t  healthy_mask;                         /* bit i: Health.Status == "Healthy" (scheduler.go:597) */
    uint8_t  allocated_mask;                       /* ENGINE-OWNED: GPUs named by ledger entries on this node; ignored on upsert */
    uint8_t  flags;                                /* bit0: record present (ENGINE-OWNED) */
    uint8_t  mig_enabled_mask;                     /* bit i: GPU i has MIGConfig != nil && Enabled (discovery.go:442) */
    uint8_t  mig_has_free_mask;                    /* bit i: ... and at least one instance with !InUse (discovery.go:444-449) */
    uint32_t n_allocs;                             /* ENGINE-OWNED: ledger entries (workloads) on this node (scheduler.go:561-566) */
    uint8_t  node_flags;                           /* KGWE_NODE_* (read by the hint scorer / exporter gauge only) */
    uint8_t  pad1[3];
} kgwe_node_image;

#define KGWE_NODE_NVSWITCH    0x01   /* NVSwitchInfo != nil && Present (prometheus_exporter.go:526) */
#define KGWE_NODE_ANY_NVLINK  0x02   /* some GPU has at least one NVLinkConnections entry, on- or off-node (prometheus_exporter.go:531-536) */

/* One pending workload: the fields of GPUWorkload.Requirements the path reads
 * (src/scheduler/types.go:36-57,80-89).  32 bytes. */
typedef struct kgwe_pod_req {
    uint64_t wl_id;               /* host-interned GPUWorkload.UID (ledger key) */
    int32_t  gpu_count;           /* Requirements.GPUCount, >= 0 */
    int32_t  min_memory_gb;       /* Requirements.MinMemoryGB, 0..2^20 */
    double   min_bandwidth_gbps;  /* EXTENSION E3 (field unused by the reference); 0 = off = reference behaviour */
    uint8_t  topology_pref;       /* KGWE_TOPO_* */
    uint8_t  arch;                /* KGWE_ARCH_ANY or the required architecture */
    uint8_t  mig_profile;         /* KGWE_MIG_NONE, 1..10, or KGWE_MIG_UNKNOWN */
    uint8_t  flags;               /* KGWE_POD_* */
    int32_t  mig_count;           /* MIGRequirements.Count */
} kgwe_pod_req;

#define KGWE_POD_EXHAUSTIVE  0x01  /* EXTENSION E1: score all C(A,k) subsets instead of the greedy (NVLink modes) */

/* Result flags */
#define KGWE_F_ELIGIBLE   0x01   /* isNodeEligible passed (Filter) */
#define KGWE_F_KEPT       0x02   /* eligible and TotalScore > 0 (scheduler.go:196) => in scoreNodes' result */
#define KGWE_F_HINT       0x04   /* hint bonus applied */
#define KGWE_F_FALLBACK   0x08   /* topology fallback arm taken (30 / 50 / 40) */
#define KGWE_F_MINBW      0x10   /* filtered by E3 min-bandwidth */
#define KGWE_F_GANG_ROLLBACK 0x20 /* E5: this member was placeable but its gang was not; nothing was committed */

/* NodeScore for one (pod, node) evaluation (src/scheduler/types.go:295-319).  16 bytes.
 * Scores are the reference's unclamped integers.  `order` is SelectedGPUs in the
 * reference's order: GPU index at position p in bits [3p, 3p+3), count in bits [24,28). */
typedef struct kgwe_node_score {
    int32_t  total;        /* TotalScore */
    int16_t  topology;     /* TopologyScore */
    int16_t  resource;     /* ResourceScore */
    int16_t  balance;      /* BalanceScore */
    uint8_t  gpu_mask;     /* SelectedGPUs as a set */
    uint8_t  flags;        /* KGWE_F_* */
    uint32_t order;        /* SelectedGPUs as an ordered list */
} kgwe_node_score;

/* A ranked placement / a SchedulingDecision (src/scheduler/types.go:253-277).  32 bytes. */
typedef struct kgwe_placement {
    uint32_t node_idx;           /* global node index, KGWE_NO_NODE when there is none */
    int32_t  total;
    int16_t  topology;
    int16_t  resource;
    int16_t  balance;
    uint8_t  gpu_mask;
    uint8_t  flags;
    uint32_t order;
    uint32_t reserved;
    double   est_bandwidth_gbps; /* EstimatedBandwidth, scheduler.go:656-692 */
} kgwe_placement;

typedef struct kgwe_stats {
    uint64_t evaluations;        /* (pod,node) pairs scored since create */
    uint64_t candidates;         /* GPU subsets whose score was actually computed (groups grown, subsets summed; 1 for an
                                  * evaluation without a group search).  Equals candidates_resolved unless a KGWE_CFG_* flag
                                  * lets evaluations share or skip parts of the search. */
    uint64_t scheduled;          /* SchedulerMetrics.TotalScheduled */
    uint64_t failed;             /* SchedulerMetrics.TotalFailed */
    uint64_t topology_optimal;   /* TopologyOptimalPlacements (TopologyScore >= 80, scheduler.go:154) */
    uint64_t kernel_launches;    /* engine kernels launched since create */
    uint64_t h2d_bytes;
    uint64_t d2h_bytes;
    double   last_call_gpu_ms;   /* device time of the last scoring call (CUDA events) */
    uint32_t n_nodes;            /* highest local index ever upserted + 1 (high-water mark; removed records score as absent) */
    uint32_t n_allocations;      /* ledger size */
    double   score_kernel_ms;    /* summed device time of the scoring kernel (K1) launches timed so far (CUDA events on the launching stream) */
    uint64_t score_kernel_timed; /* how many K1 launches that sum covers */
    uint64_t candidates_resolved;/* candidates by the reference's counting rule: A per NVLink-greedy evaluation with
                                  * k >= 2 on an eligible node, C(A,k) in exhaustive mode, 1 otherwise - whatever the engine did */
    /* Scheduling-latency histogram, the engine-side source of the reference's kgwe_scheduling_latency_ms
     * (src/monitoring/prometheus_exporter.go:258-263, read by the Grafana p99 panel, deploy/monitoring/grafana-dashboard.json:191).
     * Same bucket edges as there - 10, 25, 50, 100, 250, 500, 1000, 2500, 5000, +Inf - but in MICROSECONDS: a call here takes
     * tens of microseconds, the reference's millisecond buckets would put everything into the first one.  Counts are per
     * bucket (not cumulative).  Observed: the wall time of every kgwe_score_batch call (one observation per call), and for
     * kgwe_schedule_queue / kgwe_schedule_gangs the call's wall time divided by its workloads, once per workload. */
    uint64_t latency_us_bucket[KGWE_LATENCY_BUCKETS];
    uint64_t latency_count;
    double   latency_sum_us;
} kgwe_stats;

typedef struct kgwe_engine kgwe_engine;

/* Replaces NewTopologyAwareScheduler's state set-up (scheduler.go:97-111): device
 * image, ledger, staging.  Fails with KGWE_ERR_NO_DEVICE when CUDA is unusable. */
int kgwe_engine_create(const kgwe_config* cfg, kgwe_engine** out);
int kgwe_engine_destroy(kgwe_engine* e);

/* Fill *cfg with DefaultSchedulerConfig() values (src/scheduler/types.go:379-392). */
void kgwe_config_default(kgwe_config* cfg);

/* Topology snapshot ingestion.  Replaces the full rebuild per informer event in
 * DiscoveryService.RefreshTopology / watchNodes (src/discovery/discovery.go:290-319,
 * 589-600) by one record update; `local_idx` is the index inside this engine's shard.
 * Ledger-owned fields of *img are ignored and re-derived. */
int kgwe_node_upsert(kgwe_engine* e, uint32_t local_idx, const kgwe_node_image* img);
int kgwe_nodes_upsert(kgwe_engine* e, uint32_t first_local_idx, uint32_t count, const kgwe_node_image* imgs);
int kgwe_node_remove(kgwe_engine* e, uint32_t local_idx);
/* The same refresh for a node whose GPU LIST changed (a GPU fell off the bus, NVML enumerates in another order): the
 * ledger names GPUs by their index in the list, the reference by UUID (scheduler.go:586-598), so the host - which knows
 * the UUIDs - says where each old index went: old_to_new[i] = new index of the GPU that was at index i, -1 if it is gone
 * (KGWE_MAX_GPUS entries; values must be distinct).  Every ledger entry on the node is moved accordingly before the new
 * record is applied; an entry whose GPUs all vanished stays in the ledger and still counts as a workload on the node. */
int kgwe_node_upsert_remap(kgwe_engine* e, uint32_t local_idx, const kgwe_node_image* img, const int8_t* old_to_new);

/* Ledger.  Replaces recordAllocation (scheduler.go:695-707; same wl_id overwrites) and
 * ReleaseAllocation (scheduler.go:710-727; unknown id -> KGWE_ERR_NOT_FOUND).  A commit that names a GPU another
 * workload already holds on that node is refused with KGWE_ERR_CONFLICT and changes nothing (tryScheduleOnNode's
 * re-check, scheduler.go:634-640); a mask bit at or above the node's n_gpus is KGWE_ERR_INVALID_ARG.  Commits on an
 * index whose record is absent are kept as they are (the reference's ledger is keyed by node name and outlives nodes). */
int kgwe_alloc_commit(kgwe_engine* e, uint64_t wl_id, uint32_t local_idx, uint8_t gpu_mask);
int kgwe_alloc_release(kgwe_engine* e, uint64_t wl_id);

/* Stateless Filter + Prioritize for P workloads over every node of the image.
 * Replaces scoreNodes (scheduler.go:182-203) and everything it calls, plus the
 * descending sort at scheduler.go:144-146 in canonical order.
 *   hint_bitmap  nullable; P rows of ceil(n_nodes/8) bytes; bit n of row p set when node n
 *                is in pod p's PlacementHint.PreferredNodes (scheduler.go:270-279)
 *   out_scores   nullable; P x n_nodes records, row-major by pod
 *   topk         0..KGWE_MAX_TOPK; out_topk gets P x topk placements, best first,
 *                unused slots have node_idx == KGWE_NO_NODE
 * Host pointers; copies are done by the engine inside the call.  n_nodes here is kgwe_stats.n_nodes
 * (highest index ever upserted + 1); removed records give all-zero score records.  Calls with at
 * most 32 pods that ask only for the best placement (no matrix, topk == 1) take a single-launch,
 * copy-free path meant for per-pod Prioritize latency. */
int kgwe_score_batch(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods,
                     const uint8_t* hint_bitmap, kgwe_node_score* out_scores,
                     uint32_t topk, kgwe_placement* out_topk);

/* Same computation with every buffer already in device memory of the engine's GPU,
 * enqueued on `cuda_stream` (a cudaStream_t; NULL = the engine's own stream) without a
 * host synchronisation.  Used for device-resident pipelines (node-sharded multi-GPU
 * merge over NCCL) and for kernel-only timing. */
int kgwe_score_batch_device(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods,
                            const uint8_t* d_hint_bitmap, kgwe_node_score* d_out_scores,
                            uint32_t topk, kgwe_placement* d_out_topk, void* cuda_stream);

/* Best placement per pod WITHOUT any allocation inside the call.  No counterpart in the
 * reference: this entry point exists for hosts of this library, not for a function of KGWE.
 *
 * kgwe_score_batch_device may grow the engine's arg-max key table with cudaMalloc when a
 * batch is larger than any before it.  Some callers must never reach cudaMalloc once they
 * are serving: a scheduler extender answering Prioritize under a deadline (cudaMalloc
 * takes the driver's allocation lock and synchronises the whole device, so every other
 * engine / shard in the process stalls behind it), and processes that handed all free
 * device memory to another owner at start-up (a late cudaMalloc there can simply fail).
 * This variant computes the same result as kgwe_score_batch_device(..., topk = 1, ...)
 * with the key table in a fixed array that is part of the library image:
 * KGWE_NOALLOC_MAX_PODS keys of 8 bytes = 128 KiB, held by every GPU that loads the
 * library until the process exits.
 *
 *   - inputs and outputs are the caller's device buffers, as for kgwe_score_batch_device;
 *     d_out_scores is optional, d_out_top1 gets n_pods placements;
 *   - n_pods > KGWE_NOALLOC_MAX_PODS is refused with KGWE_ERR_TOO_LARGE (split the batch);
 *   - the call makes no cudaMalloc / cudaFree / cudaMallocHost and uses no allocating
 *     library routine; pending node upserts are still copied to the image first;
 *   - the FIRST call on a device looks up the array's device address; make that first
 *     call before starting any stream capture on that device;
 *   - all calls on one GPU share the one array (it keeps whatever the previous call left,
 *     so each call clears the slots it uses first): two calls that overlap on the same
 *     GPU - different engines or different streams - race on it.  Serialise them. */
#define KGWE_NOALLOC_MAX_PODS 16384u
int kgwe_score_top1_noalloc(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods,
                            const uint8_t* d_hint_bitmap, kgwe_node_score* d_out_scores,
                            kgwe_placement* d_out_top1, void* cuda_stream);

/* Node-sharded top-1 batches with the exchange fused into the finalisation kernel: instead of
 * writing this rank's best placement per pod locally and all-gathering, the kernel stores it
 * straight into EVERY rank's gather buffer (slot `my_rank` of a [n_ranks][n_pods] placement array)
 * through peer-mapped device memory (NVLink / NVSwitch P2P stores).  peer_gather_bufs is a HOST
 * array of n_ranks device addresses, one per rank (this rank's own buffer included), e.g. the
 * buffer_ptrs of a torch symmetric-memory allocation.  The caller still has to order "all ranks
 * have written" before reading its buffer (a cross-GPU barrier on the same stream) and then runs
 * kgwe_topk_merge_device on it.  Otherwise identical to kgwe_score_batch_device(..., topk = 1). */
int kgwe_score_batch_device_p2p(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods,
                                const uint8_t* d_hint_bitmap, kgwe_node_score* d_out_scores,
                                const uint64_t* peer_gather_bufs, uint32_t n_ranks, uint32_t my_rank, void* cuda_stream);

/* The whole node-sharded top-1 step in one call, with no collective and no separate cross-GPU barrier launch:
 * K1 on this rank's shard, the finalisation kernel stores this rank's placement of every pod into slot `my_rank` of EVERY
 * rank's gather buffer over peer-mapped memory and then raises this rank's flag on every rank, and a merge kernel waits
 * (on the device) until all ranks' flags on THIS rank have reached `epoch` before it merges into d_merged.
 *   peer_gather_bufs[r]  address, on rank r, of a buffer of 2 x n_ranks x n_pods placements (two slots, used alternately)
 *   peer_flag_bufs[r]    address, on rank r, of n_ranks + 1 uint32, zero before the first call (the last word is set
 *                        to 1 by a merge kernel that gave up waiting for a rank)
 *   epoch                1, 2, 3 ... : the same value on every rank for the same step, one higher each step
 * Host arrays of n_ranks device addresses (e.g. the buffer_ptrs of torch symmetric-memory allocations).  All ranks must
 * make the same sequence of calls.  Otherwise as kgwe_score_batch_device(..., topk = 1). */
int kgwe_score_merge_p2p(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods, const uint8_t* d_hint_bitmap,
                         kgwe_node_score* d_out_scores, const uint64_t* peer_gather_bufs, const uint64_t* peer_flag_bufs,
                         uint32_t n_ranks, uint32_t my_rank, uint32_t epoch, kgwe_placement* d_merged, void* cuda_stream);

/* Merge R ranked lists per pod (layout [r][p][topk], device memory) into one
 * [p][topk] list in canonical order (score desc, node index asc).  This is the merge
 * every rank runs after the all-gather of local top-k lists. */
int kgwe_topk_merge_device(kgwe_engine* e, const kgwe_placement* d_lists, uint32_t n_lists,
                           uint32_t n_pods, uint32_t topk, kgwe_placement* d_out, void* cuda_stream);

/* The same merge on host memory, for hosts that gather the per-shard lists themselves (gRPC,
 * gloo, one process driving several engines).  Pure host code, needs no engine and no GPU. */
int kgwe_topk_merge_host(const kgwe_placement* lists, uint32_t n_lists, uint32_t n_pods, uint32_t topk,
                         kgwe_placement* out);

/* Stateful Schedule for a queue of workloads, in order: score, pick the best node
 * (lowest index on ties), EstimatedBandwidth, commit to the ledger, next workload.
 * Replaces Schedule + tryScheduleOnNode + recordAllocation (scheduler.go:114-179,
 * 625-653, 695-707) called once per workload.  out[i].node_idx == KGWE_NO_NODE means
 * "no eligible nodes found for workload" (scheduler.go:140). */
int kgwe_schedule_queue(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods,
                        const uint8_t* hint_bitmap, kgwe_placement* out);

/* The same queue over a node-sharded cluster, on the device ("the owner rank applies the commit"; no
 * counterpart in the reference, whose scheduler is one process).  Every rank - one engine, one GPU, one shard with its
 * kgwe_config.node_base - makes this call with the SAME workloads at about the same time.  Per workload each rank's queue
 * kernel finds its shard's best node, the ranks exchange one 64-bit word (score and global node index) through peer-mapped
 * memory, and the rank that owns the overall winner commits it: out[i] holds the decision on that rank and
 * node_idx = KGWE_NO_NODE on the others (and on all of them when no node in the cluster takes the workload).
 *   peer_recv_bufs[r]  address, on rank r, of n_ranks x 4 uint64, zero before the first call (host array of n_ranks device
 *                      addresses, e.g. the buffer_ptrs of a torch symmetric-memory allocation)
 *   seq_base           number of workloads all earlier calls on these buffers have scheduled, modulo 2^24 - 1: the same on
 *                      every rank
 * Limits: 2..8 ranks, ONE RANK PER GPU (each rank's kernel waits on the device for the other ranks' kernels, which must
 * therefore be running at the same time), a shard of at most 10,240 nodes, global node indices below 2^22, workload ids not
 * yet in any rank's ledger (kgwe_alloc_release first).  A rank whose peers do not answer within about ten seconds fails
 * with KGWE_ERR_CUDA. */
int kgwe_schedule_queue_p2p(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods, const uint8_t* hint_bitmap,
                            const uint64_t* peer_recv_bufs, uint32_t n_ranks, uint32_t my_rank, uint32_t seq_base,
                            kgwe_placement* out);

/* EXTENSION E5 - gang scheduling (GangSchedulingGroup exists only as a type in the reference,
 * src/scheduler/types.go:415-444; Schedule places a workload's GPUs on ONE node, so an FSDP
 * worldSize=64 job is 8 member workloads x 8 GPUs).  Members of a gang are consecutive in
 * `pods`; gang_of[i] is the (non-decreasing) gang number of pod i.  Each gang is scheduled with
 * the kgwe_schedule_queue semantic member by member and is all-or-nothing: if any member finds
 * no node, every commit of that gang is rolled back, all its members get node_idx ==
 * KGWE_NO_NODE (members that had a node carry KGWE_F_GANG_ROLLBACK) and the queue continues
 * with the next gang.  Workload ids must be unique and not yet in the ledger. */
int kgwe_schedule_gangs(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods, const uint32_t* gang_of,
                        const uint8_t* hint_bitmap, kgwe_placement* out);

/* EXTENSION E4 - side terms of a placement (absent from the reference): for the GPU set of each
 * placement, with pair weight w(a,b) = min over both directions of the ACTIVE link bandwidth
 * (0 when either direction has none):
 *   bisection_gbps       min over all splits of the set into floor(k/2) and the rest of the summed
 *                        w across the split (35 distinct cuts at k = 8); 0 for k <= 1
 *   ring_bottleneck_gbps max over all cyclic orders of the set of the smallest w along the ring
 *                        ((k-1)! orders with the first GPU fixed); w(a,b) for k = 2; 0 for k <= 1
 * Reported only; they do not enter TotalScore.  Host buffers. */
typedef struct kgwe_side_terms {
    double bisection_gbps;
    double ring_bottleneck_gbps;
} kgwe_side_terms;
int kgwe_placement_terms(kgwe_engine* e, const kgwe_placement* placements, uint32_t n, kgwe_side_terms* out);

/* EXTENSION E2 - MIG slice bin-packing and rebalancing.  The reference only has stubs here
 * (findAvailableInstance / findGPUWithCapacity always fail, Rebalance is empty:
 * src/sharing/mig_controller.go:340-348, 407-415, 480-504); its data are the profile table
 * (src/discovery/types.go:233-239: max 7/3/2/1/1 instances of 1g/2g/3g/4g/7g) and the 10-name
 * whitelist.  Own spec:
 *   - a GPU has 8 memory slices; `occ` is the occupancy bitmask.  A profile (ids 1..10, whitelist
 *     order) has a class (odd id = 40 GB parts: 1g.5gb 2g.10gb 3g.20gb 4g.20gb 7g.40gb; even id =
 *     80 GB parts) and a geometry with fixed legal placements (NVIDIA's MIG placement rules):
 *     1g: 1 slice at 0..6; 2g: 2 slices at 0,2,4; 3g: 4 slices at 0,4; 4g: 4 slices at 0; 7g: all 8.
 *   - a request fits placement m on a MIG-enabled GPU of its class when (occ & m) == 0.  Among all
 *     fits in the cluster choose the one that destroys the fewest still-free legal placements of
 *     that GPU (the 14 placements above, itself included); ties -> lowest GPU index, then lowest
 *     start.  Requests are served in order, each commit (occ |= m) is seen by the next. */
typedef struct kgwe_mig_gpu { uint8_t occ; uint8_t cls; uint8_t enabled; uint8_t pad; } kgwe_mig_gpu;
typedef struct kgwe_mig_assignment {
    uint32_t gpu_idx;      /* KGWE_NO_NODE when nothing fits */
    uint8_t  start, size, cost, profile;
} kgwe_mig_assignment;
/* gpus: in/out (occupancy after all commits); profiles: n_reqs profile ids; host buffers. */
int kgwe_mig_pack(kgwe_engine* e, kgwe_mig_gpu* gpus, uint32_t n_gpus, const uint8_t* profiles, uint32_t n_reqs,
                  kgwe_mig_assignment* out);

/*   - rebalancing pass: idle GPUs (enabled, occ == 0) of each class are split among that class's
 *     profiles in proportion to target[profile-1] (MIGStrategy.ProfileDistribution, mig_controller.go
 *     :263-270): profile p gets floor(share_p * idle) GPUs in profile order, the remainder goes to the
 *     largest fractional parts (ties: lowest profile id); idle GPUs are handed out in index order and
 *     each is filled with as many instances of its profile as fit (7/3/2/1/1).  Emits one create op
 *     per instance - 40 GB-class GPUs first, then 80 GB-class, each by profile, GPU index, start -
 *     and the resulting free-instance supply per profile.  *n_ops is the number of ops the plan
 *     has; only the first ops_cap are written. */
typedef struct kgwe_mig_op { uint32_t gpu_idx; uint8_t profile, start, size, pad; } kgwe_mig_op;      /* pad: KGWE_MIG_OP_* (always CREATE from kgwe_mig_rebalance) */
int kgwe_mig_rebalance(kgwe_engine* e, const kgwe_mig_gpu* gpus, uint32_t n_gpus, const double* target /*[10]*/,
                       kgwe_mig_op* ops, uint32_t ops_cap, uint32_t* n_ops, uint32_t* supply /*[10]*/);

/*   - reconfiguration pass (the reference sketches it: Rebalance compares usage against strategy and reconfigures when
 *     AllowDynamicReconfig is set, src/sharing/mig_controller.go:480-504 - its body is empty).  Here a GPU is described by
 *     the instances that EXIST on it and the ones IN USE, as bit sets over the 14 legal placements
 *     (bits 0..6: 1g at slice 0..6; 7..9: 2g at 0,2,4; 10..11: 3g at 0,4; 12: 4g at 0; 13: 7g), in_use a subset of created.
 *     A GPU is reconfigurable when it is enabled and nothing on it is in use; running instances are never touched.
 *     The reconfigurable GPUs of each class are split among that class's profiles by the same rule as kgwe_mig_rebalance
 *     (shares of target[], largest remainders, GPUs in index order).  A GPU given geometry q ends up with exactly the
 *     instances of q at all its legal starts: for it the pass emits a DESTROY op for every existing instance that is not
 *     one of those (ascending placement bit) and then a CREATE op for every missing one (ascending start) - a GPU that
 *     already has the wanted layout emits nothing.  Ops come out by class, profile, GPU index.  supply[p] counts the free
 *     instances of profile p after the pass (reconfigured GPUs, plus the existing free instances of GPUs left alone). */
typedef struct kgwe_mig_gpu_state { uint16_t created; uint16_t in_use; uint8_t cls; uint8_t enabled; uint8_t pad[2]; } kgwe_mig_gpu_state;
#define KGWE_MIG_OP_CREATE  0      /* kgwe_mig_op.pad: what the op does */
#define KGWE_MIG_OP_DESTROY 1
int kgwe_mig_reconfigure(kgwe_engine* e, const kgwe_mig_gpu_state* gpus, uint32_t n_gpus, const double* target /*[10]*/,
                         kgwe_mig_op* ops, uint32_t ops_cap, uint32_t* n_ops, uint32_t* supply /*[10]*/);

/* DiscoveryService.GetTopologyHint (src/discovery/discovery.go:222-247) - the reference's second
 * node-parallel scorer, over the same image ("next" row 8f.3).  Per request: every node is scored
 * by scoreNodeForRequirements (discovery.go:378-434: availability rule :436-460 - healthy and
 * either a free MIG instance or GPUPercent < 90 -, free-memory filter, base 50, +30 when
 * findNVLinkGroups (:462-488) finds a GPU with Count-1 linked peers among the filtered GPUs, +10
 * same NUMA, +5 per selected GPU of the wanted architecture) and the best node wins
 * (score > best, :231).  Canonical order: lowest node index on ties, peers of a GPU in ascending
 * GPU index (Go iterates a map there).  out[i].node_idx == KGWE_NO_NODE <=> "no suitable placement
 * found for requirements" (:241).  The ledger is not consulted (the reference does not either). */
typedef struct kgwe_hint_req {
    int32_t count;             /* GPURequirements.Count */
    int32_t min_memory_gb;     /* MinMemoryGB */
    uint8_t prefer_nvlink;     /* PreferNVLink */
    uint8_t prefer_same_numa;  /* PreferSameNUMA */
    uint8_t arch;              /* Architecture, KGWE_ARCH_ANY = "" */
    uint8_t pad[5];
} kgwe_hint_req;
typedef struct kgwe_topology_hint {
    uint32_t node_idx;         /* PreferredNode */
    int32_t  score;            /* Score */
    uint32_t order;            /* PreferredGPUs, packed like kgwe_node_score.order */
    uint8_t  gpu_mask;
    uint8_t  pad[3];
    double   bandwidth_estimate;   /* BandwidthEstimate, discovery.go:506-541 */
} kgwe_topology_hint;
int kgwe_topology_hint_batch(kgwe_engine* e, const kgwe_hint_req* reqs, uint32_t n_reqs, kgwe_topology_hint* out);

/* PrometheusExporter.calculateTopologyScore per node (src/monitoring/prometheus_exporter.go:517-539,
 * "next" row 8f.4): 0 without GPUs, else min(100, 50 + 30 NVSwitch + 20 any NVLink entry).
 * out gets one int32 per record 0..n_nodes-1 (absent records: 0). */
int kgwe_topology_gauges(kgwe_engine* e, int32_t* out, uint32_t n_nodes);

/* Checkpoint / resume of the engine state (cluster image + ledger).  The reference keeps its ledger
 * in memory only (scheduler.go:29) and rebuilds topology from NVML/Kubernetes; a scheduler that
 * restarts can use these to come back without replaying every event.  The blob is opaque,
 * versioned and self-describing; kgwe_snapshot_load replaces image and ledger of an engine whose
 * max_nodes is large enough (weights / normalisation stay those of the loading engine). */
int kgwe_snapshot_size(kgwe_engine* e, uint64_t* bytes);
int kgwe_snapshot_save(kgwe_engine* e, void* buf, uint64_t capacity, uint64_t* written);
int kgwe_snapshot_load(kgwe_engine* e, const void* buf, uint64_t len);

/* Read one record back (device -> host) in ABI layout, ledger fields included. */
int kgwe_node_read(kgwe_engine* e, uint32_t local_idx, kgwe_node_image* out);

/* Counters so far.  Does not synchronise the device: kernels of asynchronous (device-pointer) calls that are still
 * running show up in a later poll, and so do the K1 timings of launches that have not finished. */
int kgwe_get_stats(kgwe_engine* e, kgwe_stats* out);
const char* kgwe_last_error(kgwe_engine* e);
const char* kgwe_version(void);

#ifdef __cplusplus
}
#endif
#endif /* KGWE_ENGINE_H */
