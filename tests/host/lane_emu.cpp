// lane_emu.cpp — TEST INFRASTRUCTURE: K1's per-lane evaluation (csrc/kgwe_lane.cuh) compiled for the host.
//
// The scoring kernel gives every lane one node and runs plain scalar code per (pod, node) evaluation, so that code
// can also be built with g++ once the CUDA intrinsics it uses have host stand-ins.  This file does that and
// exposes one C function that scores a pod batch node by node exactly the way a lane does (same set-up functions,
// same tables, laid out with stride 1).  tests/test_lane_emu.py compares its output with the oracle, which checks
// the kernel's logic on machines without a GPU.  It is not part of the engine: nothing under
// k8s-gpu-workload-enhancer_b200/ builds, links or loads it, and the engine has no CPU path.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>      // vector types only (int4, double2, make_int4 ...); __device__ expands to nothing under g++

// ---- host stand-ins for the device intrinsics used by kgwe_device.cuh / kgwe_lane.cuh ----
#define __noinline__ __attribute__((noinline))
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
static inline unsigned __brev(unsigned x) { unsigned r = 0; for (int i = 0; i < 32; ++i) if (x >> i & 1u) r |= 1u << (31 - i); return r; }
static inline unsigned __byte_perm(unsigned x, unsigned y, unsigned s) {
    const uint64_t src = ((uint64_t)y << 32) | x;
    unsigned r = 0;
    for (int i = 0; i < 4; ++i) {
        const unsigned sel = (s >> (4 * i)) & 0xf;
        unsigned b = (unsigned)(src >> (8 * (sel & 7))) & 0xff;
        if (sel & 8) b = (b & 0x80) ? 0xff : 0x00;
        r |= b << (8 * i);
    }
    return r;
}
static inline double __dadd_rn(double a, double b) { return a + b; }      // built with -ffp-contract=off
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __ddiv_rn(double a, double b) { return a / b; }
static inline int __double2int_rz(double x) { return x >= 2147483647.0 ? 2147483647 : x <= -2147483648.0 ? (-2147483647 - 1) : (x != x ? 0 : (int)x); }
static inline long long __double_as_longlong(double x) { long long r; memcpy(&r, &x, 8); return r; }
static inline double __longlong_as_double(long long x) { double r; memcpy(&r, &x, 8); return r; }
static inline int __vimax3_s32(int a, int b, int c) { return std::max(a, std::max(b, c)); }
using std::max;
using std::min;

#include "../../k8s-gpu-workload-enhancer_b200/csrc/kgwe_lane.cuh"
#include "../../k8s-gpu-workload-enhancer_b200/csrc/kgwe_queue_state.cuh"
#include "../../k8s-gpu-workload-enhancer_b200/csrc/kgwe_pack.hpp"
#include "../../oracle/kgwe_oracle.h"

using namespace kgwe;

extern "C" int lane_emu_score_batch(const kgwe_config* cfg, const kgwe_node_image* nodes, uint32_t n_nodes,
                                    const kgwe_oracle_alloc* ledger, uint32_t n_ledger,
                                    const kgwe_pod_req* pods, uint32_t n_pods, const uint8_t* hint,
                                    kgwe_node_score* out, int general, int tables, uint32_t mode,
                                    uint64_t* scored_out, uint64_t* resolved_out) {
    std::vector<DevNode> img(n_nodes);
    char msg[160];
    for (uint32_t n = 0; n < n_nodes; ++n) {
        if (to_device_record(nodes[n], img[n], msg, sizeof msg) != 0) return -1;
        if (!(nodes[n].flags & 1)) memset(&img[n], 0, sizeof(DevNode));      // absent record
    }
    for (uint32_t i = 0; i < n_ledger; ++i) {
        if (ledger[i].node_idx >= n_nodes) return -2;
        img[ledger[i].node_idx].allocated_mask |= ledger[i].gpu_mask;
        img[ledger[i].node_idx].n_allocs++;
    }
    const Params P{cfg->topology_weight, cfg->resource_weight, cfg->balance_weight, cfg->hint_bonus, cfg->nvlink_norm_gbps, cfg->node_base};
    const uint32_t row_bytes = (n_nodes + 7) / 8;
    uint16_t slot_of[256];
    for (int m = 0; m < 256; ++m) {
        int rank = 0;
        for (int q = 0; q < m; ++q) rank += __popc(q) == __popc(m);
        slot_of[m] = (uint16_t)(subset_off(__popc(m)) + rank);
    }
    std::vector<PodLite> pl(n_pods);
    for (uint32_t p = 0; p < n_pods; ++p) pl[p] = pod_lite(pods[p]);
    uint64_t scored_sum = 0, resolved_sum = 0;
    for (uint32_t n = 0; n < n_nodes; ++n) {
        const DevNode* rec = &img[n];
        NodeCtx c;
        node_ctx_core(c, rec, true);
        node_ctx_group_tables(c, rec);
        LaneNode ln;
        lane_node_init(ln, c, rec, general == 0 ? 0u : (uint32_t)general);
        int32_t topo[7] = {0}; uint32_t grp[6] = {0}, mig[3] = {0}; int2 memo[7]; int4 mat[32]; std::vector<uint32_t> subset(SUBSET_ENTRIES, 0u);
        memset(memo, 0, sizeof memo); memset(mat, 0, sizeof mat);
        LaneTabs tb{topo, grp, mig, memo, mat, subset.data(), 1};
        if (tables) lane_tables_build<true, true>(ln, c, tb, rec, P.norm, mode, slot_of);
        else if (general) lane_tables_build<true, false>(ln, c, tb, rec, P.norm, mode, slot_of);
        else lane_tables_build<false, false>(ln, c, tb, rec, P.norm, mode, slot_of);
        for (uint32_t p = 0; p < n_pods; ++p) {
            const bool hinted = hint && ((hint[(size_t)p * row_bytes + (n >> 3)] >> (n & 7)) & 1);
            LaneResult r;
            const double min_bw = pods[p].min_bandwidth_gbps;
            if (tables) lane_evaluate<true, true>(ln, tb, rec, pl[p], min_bw, hinted, P, mode, r);
            else if (general) lane_evaluate<true, false>(ln, tb, rec, pl[p], min_bw, hinted, P, mode, r);
            else lane_evaluate<false, false>(ln, tb, rec, pl[p], min_bw, hinted, P, mode, r);
            const uint32_t w[4] = {r.w0, r.w1, r.w2, r.w3};
            memcpy(&out[(size_t)p * n_nodes + n], w, 16);
            // the arg-max contribution must be TotalScore exactly for kept evaluations
            const kgwe_node_score& rec16 = out[(size_t)p * n_nodes + n];
            if (((rec16.flags & KGWE_F_KEPT) ? (uint32_t)rec16.total : 0u) != r.kept_total) return -3;
            scored_sum += r.scored; resolved_sum += r.candidates;
        }
    }
    if (scored_out) *scored_out = scored_sum;
    if (resolved_out) *resolved_out = resolved_sum;
    return 0;
}

// The cluster queue kernel's owner refresh (qc_state on the static columns, kgwe_queue_state.cuh) against the set-up path
// every other kernel uses (node_ctx_core + lane_node_init on the record): for every node and every (allocated_mask,
// n_allocs) in `allocs` (pairs of mask, count) all LaneNode fields must agree.  Returns the number of disagreements
// (-1: a record was refused by the packer).
extern "C" long long lane_emu_queue_state_check(const kgwe_node_image* nodes, uint32_t n_nodes, const uint32_t* allocs, uint32_t n_allocs_cases) {
    qc_first_tables(qc_prefix_lut, qc_order_lut);
    long long bad = 0;
    char msg[160];
    for (uint32_t n = 0; n < n_nodes; ++n) {
        DevNode rec;
        if (to_device_record(nodes[n], rec, msg, sizeof msg) != 0) return -1;
        const bool present = nodes[n].flags & 1;
        if (!present) memset(&rec, 0, sizeof rec);
        NodeCtx c0; node_ctx_core(c0, &rec, true);
        LaneNode l0; lane_node_init(l0, c0, &rec, 0u);
        uint64_t st[QS_ENTRIES];
        qc_static_build(st, 1, &rec, c0, l0.tier);
        for (uint32_t i = 0; i < n_allocs_cases; ++i) {
            DevNode r2 = rec;
            if (present) { r2.allocated_mask = (uint8_t)allocs[2 * i]; r2.n_allocs = allocs[2 * i + 1]; }
            NodeCtx c; node_ctx_core(c, &r2, true);
            LaneNode want; lane_node_init(want, c, &r2, 0u);
            LaneNode got; memset(&got, 0, sizeof got);
            qc_state(got, st, 1, present ? r2.allocated_mask : 0u, present ? r2.n_allocs : 0u);
            const bool same = got.avail == want.avail && got.A == want.A && got.rlo == want.rlo && got.rhi == want.rhi && got.sum_free == want.sum_free
                           && got.util_bonus == want.util_bonus && got.balance == want.balance && got.arch_lo == want.arch_lo && got.arch_hi == want.arch_hi
                           && got.avail_order == want.avail_order && got.avail_prefix == want.avail_prefix && got.tier == want.tier
                           && got.valid == want.valid && got.integral == want.integral;
            bad += same ? 0 : 1;
        }
    }
    return bad;
}
