// kgwe_queue_state.cuh — the cluster queue kernel's per-lane node state (kgwe_queue.cuh): the static shared-memory
// columns of a node and the derivation of the lane's pod-independent state from them and the node's current
// allocations.  Plain scalar code like kgwe_lane.cuh, so the test suite also compiles it for the host
// (tests/host/lane_emu.cpp) and checks it against node_ctx_core + lane_node_init for every allocation state.
#pragma once
#include "kgwe_lane.cuh"

namespace kgwe {

// Per-lane columns of 8-byte entries that a commit never changes (entry e of a lane at st[e * stride]).
constexpr int QS_UTIL = 0;       // [8] Utilization.GPUPercent per GPU (double)
constexpr int QS_FREE = 8;       // [8] Memory.FreeBytes per GPU
constexpr int QS_ROWS = 16;      // adjacency rows of the greedy graph, reversed bit order (LaneNode::rlo/rhi before the restriction to available GPUs)
constexpr int QS_ARCH = 17;      // arch[0..7]
constexpr int QS_NUMA = 18;      // numa[0..7]
constexpr int QS_PCIE = 19;      // pcie_switch[0..7]
constexpr int QS_BAL = 20;       // BalanceScore for 0..7 allocations on the node, one byte each (0 from n_gpus allocations on)
constexpr int QS_MISC = 21;      // healthy_mask | n_gpus (<= 8) << 8 | tier << 16 | valid << 24 | integral << 25
constexpr int QS_ENTRIES = 22;

// avail (8-bit set) -> avail_prefix (byte i: the first i + 1 members as a set) and avail_order (members in index order, 3 bits each):
// NodeCtx::avail_prefix / avail_order (node_ctx_core) as a table, filled by the engine at start-up (qc_first_tables).
#ifdef __CUDACC__
__constant__ uint64_t qc_prefix_lut[256];
__constant__ uint32_t qc_order_lut[256];
#else
static uint64_t qc_prefix_lut[256];      // host-side emulation: filled by the test driver
static uint32_t qc_order_lut[256];
#endif
inline void qc_first_tables(uint64_t* prefix, uint32_t* order) {
    for (uint32_t avail = 0; avail < 256; ++avail) {
        uint64_t pf = 0; uint32_t od = 0, acc = 0; int i = 0;
        for (uint32_t b = 0; b < 8; ++b)
            if (avail >> b & 1u) { acc |= 1u << b; od |= b << (3 * i); pf |= (uint64_t)acc << (8 * i); ++i; }
        prefix[avail] = pf; order[avail] = od;
    }
}

// bit i of the low nibble of x -> byte i all ones
__device__ __forceinline__ uint32_t qc_expand4(uint32_t x) { return (((x & 0xfu) * 0x00204081u) & 0x01010101u) * 0xffu; }

// The static columns of this lane's node (once, at set-up).
__device__ inline void qc_static_build(uint64_t* st, int S, const DevNode* rec, const NodeCtx& c, uint32_t tier) {
    uint32_t ng = c.valid ? rec->n_gpus : 0; ng = ng > 8 ? 8 : ng;
    uint64_t rows = 0, bal = 0;
    for (int g = 0; g < 8; ++g) {
        st[(QS_UTIL + g) * S] = (uint64_t)__double_as_longlong(c.valid ? rec->util[g] : 0.0);
        st[(QS_FREE + g) * S] = c.valid ? rec->free_bytes[g] : 0ull;
        rows |= (uint64_t)rev8((uint32_t)(c.colact >> (8 * g)) & 0xffu) << (8 * (7 - g));
        if (ng > 0) {       // calculateBalanceScore (scheduler.go:556-578) with g allocations on the node
            const double usage = __ddiv_rn((double)g, (double)ng);
            const int s = __double2int_rz(__dmul_rn(100.0, __dsub_rn(1.0, usage)));
            bal |= (uint64_t)(s > 0 ? s : 0) << (8 * g);
        }
    }
    st[QS_ROWS * S] = rows; st[QS_ARCH * S] = c.arch8; st[QS_BAL * S] = bal;
    st[QS_NUMA * S] = c.valid ? ld_u64(rec->numa) : 0ull; st[QS_PCIE * S] = c.valid ? ld_u64(rec->pcie_switch) : 0ull;
    st[QS_MISC * S] = (uint64_t)((c.valid ? (uint32_t)rec->healthy_mask : 0u) | (ng << 8) | (tier << 16) | ((c.valid ? 1u : 0u) << 24) | ((c.integral ? 1u : 0u) << 25));
}

// The lane's pod-independent state (node_ctx_core + lane_node_init) from the static columns and the node's current
// allocations: what the owner of a winning node runs after its commit, on the critical path of the queue, so it
// reads shared memory only and divides nothing.  The utilisation bonus compares the ordered sum with 30 * A and
// 60 * A instead of the mean with 30 and 60: for A <= 8 both thresholds times A are exact doubles whose spacing
// exceeds A half-ulps of the threshold, so no sum below the product has a quotient that rounds up to the threshold.
__device__ __forceinline__ void qc_state(LaneNode& ln, const uint64_t* st, int S, uint32_t allocated, uint32_t n_allocs) {
    const uint32_t misc = (uint32_t)st[QS_MISC * S];
    const uint32_t ng = (misc >> 8) & 0xffu;
    ln.valid = (misc >> 24) & 1u; ln.integral = (misc >> 25) & 1u; ln.tier = (misc >> 16) & 0xffu;
    const uint32_t avail = (misc & 0xffu) & ~allocated & ((1u << ng) - 1u);       // misc = 0 for an absent record
    ln.avail = avail; ln.A = __popc(avail);
    uint64_t sum = 0; double usum = 0.0;
#pragma unroll
    for (int g = 0; g < 8; ++g)
        if (avail >> g & 1) { sum += st[(QS_FREE + g) * S]; usum = __dadd_rn(usum, __longlong_as_double((long long)st[(QS_UTIL + g) * S])); }
    ln.sum_free = sum;
    const double a = (double)ln.A;
    ln.util_bonus = ln.A > 0 ? (usum < __dmul_rn(30.0, a) ? 25 : (usum < __dmul_rn(60.0, a) ? 10 : 0)) : 0;
    ln.balance = n_allocs < 8u ? (int)((st[QS_BAL * S] >> (8 * n_allocs)) & 0xffu) : 0;
    const uint32_t availR = rev8(avail);
    const uint64_t rows = st[QS_ROWS * S], arch = st[QS_ARCH * S];
    ln.rlo = (uint32_t)rows & (availR * 0x01010101u) & qc_expand4(availR);
    ln.rhi = (uint32_t)(rows >> 32) & (availR * 0x01010101u) & qc_expand4(availR >> 4);
    ln.arch_lo = (uint32_t)arch & qc_expand4(avail);
    ln.arch_hi = (uint32_t)(arch >> 32) & qc_expand4(avail >> 4);
    ln.avail_prefix = qc_prefix_lut[avail]; ln.avail_order = qc_order_lut[avail];
}

}  // namespace kgwe
