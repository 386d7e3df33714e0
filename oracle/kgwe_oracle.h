/*
 * kgwe_oracle.h — C API of the CPU oracle.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker or the timed CPU baseline —
 * never as part of the product path (the CUDA engine has no CPU fallback).
 *
 * PARITY UNPINNED BY REFERENCE TESTS: the reference ships no tests, golden vectors or
 * fixtures for this path and its Go toolchain is absent here, so this restatement is
 * pinned by source reading plus the hand-derived known-answer vectors of SURVEY.md 8c
 * (tests/golden/kats.json), not by outputs of the reference itself.
 */
#ifndef KGWE_ORACLE_H
#define KGWE_ORACLE_H

#include "../include/kgwe_engine.h"

#ifdef __cplusplus
extern "C" {
#endif

/* A ledger entry (GPUAllocation, src/scheduler/scheduler.go:68-75) in index form. */
typedef struct kgwe_oracle_alloc {
    uint64_t wl_id;
    uint32_t node_idx;
    uint8_t  gpu_mask;
    uint8_t  pad[3];
} kgwe_oracle_alloc;

enum { KGWE_ORACLE_FAITHFUL = 0, KGWE_ORACLE_INDEXED = 1 };

typedef struct kgwe_oracle_counters {
    uint64_t evaluations;
    uint64_t candidates;
    double   seconds;          /* wall time of the scoring work only (structures already built) */
} kgwe_oracle_counters;

/* scoreNodes for each pod over all nodes (stateless w.r.t. the ledger passed in).
 * mode FAITHFUL keeps the reference's data structures (string UUIDs, hash-map ledger
 * scanned per node, per-call adjacency maps); INDEXED computes the same results on
 * dense arrays.  `threads` > 1 splits pods over std::threads. */
int kgwe_oracle_score_batch(const kgwe_config* cfg,
                            const kgwe_node_image* nodes, uint32_t n_nodes,
                            const kgwe_oracle_alloc* ledger, uint32_t n_ledger,
                            const kgwe_pod_req* pods, uint32_t n_pods,
                            const uint8_t* hint_bitmap,
                            kgwe_node_score* out_scores,
                            uint32_t topk, kgwe_placement* out_topk,
                            int mode, int threads, kgwe_oracle_counters* counters);

/* Schedule() for each pod in order, committing as it goes.  ledger_out (capacity
 * n_ledger + n_pods) receives the final ledger; *n_ledger_out its size. */
int kgwe_oracle_schedule_queue(const kgwe_config* cfg,
                               const kgwe_node_image* nodes, uint32_t n_nodes,
                               const kgwe_oracle_alloc* ledger, uint32_t n_ledger,
                               const kgwe_pod_req* pods, uint32_t n_pods,
                               const uint8_t* hint_bitmap,
                               kgwe_placement* out,
                               kgwe_oracle_alloc* ledger_out, uint32_t* n_ledger_out,
                               int mode, kgwe_oracle_counters* counters);

/* EXTENSION E5: all-or-nothing gangs over Schedule() (see kgwe_schedule_gangs). */
int kgwe_oracle_schedule_gangs(const kgwe_config* cfg,
                               const kgwe_node_image* nodes, uint32_t n_nodes,
                               const kgwe_oracle_alloc* ledger, uint32_t n_ledger,
                               const kgwe_pod_req* pods, uint32_t n_pods, const uint32_t* gang_of,
                               const uint8_t* hint_bitmap,
                               kgwe_placement* out,
                               kgwe_oracle_alloc* ledger_out, uint32_t* n_ledger_out, int mode);

/* EXTENSION E4: bisection bandwidth / ring bottleneck of each placement (see kgwe_placement_terms).
 * Plain loops: every subset, every permutation (std::next_permutation). */
int kgwe_oracle_placement_terms(const kgwe_config* cfg, const kgwe_node_image* nodes, uint32_t n_nodes,
                                const kgwe_placement* placements, uint32_t n, kgwe_side_terms* out);

/* DiscoveryService.GetTopologyHint (src/discovery/discovery.go:222-247, 378-541) and the exporter's
 * per-node topology gauge (src/monitoring/prometheus_exporter.go:517-539), restated on the image. */
int kgwe_oracle_topology_hints(const kgwe_config* cfg, const kgwe_node_image* nodes, uint32_t n_nodes,
                               const kgwe_hint_req* reqs, uint32_t n_reqs, kgwe_topology_hint* out);
int kgwe_oracle_topology_gauges(const kgwe_node_image* nodes, uint32_t n_nodes, int32_t* out);

/* EXTENSION E2: MIG bin-packing and rebalancing by plain loops (spec in include/kgwe_engine.h). */
int kgwe_oracle_mig_pack(kgwe_mig_gpu* gpus, uint32_t n_gpus, const uint8_t* profiles, uint32_t n_reqs, kgwe_mig_assignment* out);
int kgwe_oracle_mig_rebalance(const kgwe_mig_gpu* gpus, uint32_t n_gpus, const double* target,
                              kgwe_mig_op* ops, uint32_t ops_cap, uint32_t* n_ops, uint32_t* supply);
int kgwe_oracle_mig_reconfigure(const kgwe_mig_gpu_state* gpus, uint32_t n_gpus, const double* target,
                                kgwe_mig_op* ops, uint32_t ops_cap, uint32_t* n_ops, uint32_t* supply);

#ifdef __cplusplus
}
#endif
#endif
