// kgwe_lane.cuh — the scoring kernel's (K1) evaluation: ONE LANE = ONE NODE.
//
// A lane keeps its node's pod-independent state in registers for the whole pod chunk (availability, the
// adjacency rows of the greedy graph, resource sums, balance), the pod descriptors are decoded once per CTA
// into 16-byte PodLite records in shared memory, and an evaluation is straight-line scalar code with no
// cross-lane traffic.  The pod is the same for all 32 lanes of a warp, so every branch on a pod field
// (GPUCount, preference, architecture, MIG, minimum bandwidth) is warp-uniform: no divergence, no bucketing
// of the batch needed.  Per-node tables too large for registers live in shared memory with the LANE as the
// fastest index ([entry][lane]), which makes every access conflict-free whatever entry a lane picks.
//
// Group search of findBestNVLinkGroup (reference src/scheduler/scheduler.go:376-435), per lane, in the
// cheapest exact tier its node allows:
//   uniform   every ACTIVE link of the node carries the same bandwidth: all candidates still linked to all
//             members have the same aggregate, so "largest aggregate, first candidate on ties" is "lowest-
//             numbered candidate still linked to all members" - a walk over 8-bit sets.  The sets are kept
//             in REVERSED bit order (GPU g at bit 7 - g) so that "lowest-numbered" is "highest bit":
//             one FLO (bfind) finds it, one PRMT fetches its adjacency row out of the 64-bit row word, one
//             AND shrinks the set, one IMAD appends it to the join order.  All eight starts are walked for
//             every evaluation (one candidate group per available start GPU, :389), fully unrolled.
//   int32     whole-number bandwidths: 64 * bw in an int32 matrix ([row m][half][lane] 16-byte units),
//             running sums seeded with the tie-break code, VIMNMX3 trees (see NodeFast in kgwe_device.cuh).
//   float64   anything else: the reference's arithmetic verbatim, rows from the lane's column.
// All functions here are plain scalar code: the test suite also compiles them for the host (tests/host/
// lane_emu.cpp, with stand-ins for the intrinsics) to check the logic against the oracle without a GPU;
// that build is test infrastructure and is not linked into the engine.
#pragma once
#include "kgwe_device.cuh"

namespace kgwe {

// ---- pod descriptor, decoded once per CTA ---------------------------------------------------------
struct __align__(16) PodLite {
    uint64_t required;     // MinMemoryGB * 2^30 * k (scheduler.go:533), with the wrap-around of the 64-bit form
    int32_t  mig_count;    // MIGRequirements.Count
    uint32_t pk;           // k | cls << 8 | arch << 16 | mig_profile << 24
};
static_assert(sizeof(PodLite) == 16, "PodLite is one 16-byte shared-memory word");
// k = GPUCount clamped to 0..9; PL_K_NEVER for a negative count (never eligible).  cls bits (already shifted):
constexpr uint32_t PL_NVLINK = 0x0100;   // NVLink preference with k >= 2: the group search runs
constexpr uint32_t PL_NUMA   = 0x0200;   // SameNUMA with k >= 2
constexpr uint32_t PL_PCIE   = 0x0400;   // SamePCIeSwitch with k >= 2
constexpr uint32_t PL_EXH    = 0x0800;   // E1 exhaustive subset mode (implies PL_NVLINK)
constexpr uint32_t PL_MINBW  = 0x1000;   // E3 minimum bandwidth > 0 with k >= 2
constexpr uint32_t PL_ARCH   = 0x2000;
constexpr uint32_t PL_MIG    = 0x4000;
constexpr uint32_t PL_K_NEVER = 15;

__device__ inline PodLite pod_lite(const kgwe_pod_req& pod) {
    PodLite pl;
    const int k = clamp_k(pod.gpu_count);
    pl.required = (uint64_t)(uint32_t)pod.min_memory_gb * 1073741824ull * (uint64_t)(uint32_t)k;
    pl.mig_count = pod.mig_count;
    const bool nv = k >= 2 && is_nvlink_pref(pod);
    uint32_t c = k < 0 ? PL_K_NEVER : (uint32_t)k;
    if (nv) c |= PL_NVLINK;
    if (nv && (pod.flags & KGWE_POD_EXHAUSTIVE)) c |= PL_EXH;
    if (k >= 2 && pod.topology_pref == KGWE_TOPO_SAME_NUMA) c |= PL_NUMA;
    if (k >= 2 && pod.topology_pref == KGWE_TOPO_SAME_PCIE_SWITCH) c |= PL_PCIE;
    if (k >= 2 && pod.min_bandwidth_gbps > 0.0) c |= PL_MINBW;
    if (pod.arch != KGWE_ARCH_ANY) c |= PL_ARCH;
    if (pod.mig_profile != KGWE_MIG_NONE) c |= PL_MIG;
    pl.pk = c | ((uint32_t)pod.arch << 16) | ((uint32_t)pod.mig_profile << 24);
    return pl;
}

// ---- per-lane node state --------------------------------------------------------------------------
constexpr uint32_t TIER_NONE = 0;         // absent record / index past the image
constexpr uint32_t TIER_UNIFORM = 1;      // bit-set walks
constexpr uint32_t TIER_INT = 2;          // int32 chains on the lane's matrix column
constexpr uint32_t TIER_F64 = 3;          // float64 chains on the lane's matrix column
constexpr uint32_t TIER_F64_GLOBAL = 4;   // float64 chains reading the record itself (kernel variant without columns)

struct LaneNode {
    uint32_t avail;          // healthy & ~allocated & n_gpus bits (GPU g at bit g)          scheduler.go:581-603
    int      A;
    uint32_t rlo, rhi;       // byte b: GPUs (reversed bits) with an ACTIVE link to GPU 7-b, both available; 0 if GPU 7-b is not
    uint64_t sum_free;       // sum of FreeBytes over available GPUs                          scheduler.go:528-531
    int      util_bonus, balance;
    uint32_t arch_lo, arch_hi;   // arch[] of the available GPUs, 0 elsewhere
    uint32_t avail_order; uint64_t avail_prefix;
    uint32_t tier;
    bool     valid, integral;
};

__device__ __forceinline__ uint32_t rev8(uint32_t x) { return __brev(x) >> 24; }

// mat_units: 16-byte units of the lane's matrix column - 0 (no column), 16 (int32 rows only), 32 (int32 or float64 rows)
__device__ inline void lane_node_init(LaneNode& ln, const NodeCtx& c, const DevNode* rec, uint32_t mat_units) {
    ln.avail = c.avail; ln.A = c.A; ln.sum_free = c.sum_free; ln.util_bonus = c.util_bonus; ln.balance = c.balance;
    ln.avail_order = c.avail_order; ln.avail_prefix = c.avail_prefix; ln.valid = c.valid; ln.integral = c.integral;
    uint64_t a8 = c.arch8, rows = 0;
    for (int g = 0; g < 8; ++g) {
        if (!(c.avail >> g & 1)) { a8 &= ~(0xffull << (8 * g)); continue; }
        const uint32_t row = (uint32_t)(c.colact >> (8 * g)) & 0xffu & c.avail;      // c's with an active link to g (never g itself)
        rows |= (uint64_t)rev8(row) << (8 * (7 - g));
    }
    ln.arch_lo = (uint32_t)a8; ln.arch_hi = (uint32_t)(a8 >> 32);
    ln.rlo = (uint32_t)rows; ln.rhi = (uint32_t)(rows >> 32);
    ln.tier = !c.valid ? TIER_NONE : node_uniform_code(rec) ? TIER_UNIFORM : (c.integral && mat_units >= 16) ? TIER_INT : (!c.integral && mat_units >= 32) ? TIER_F64 : TIER_F64_GLOBAL;
}

// Per-lane views of the shared-memory tables: entry e of table X is X[e * stride] (stride = lanes per CTA;
// 1 in the host-side emulation).  Null when the kernel variant has no such table.
struct LaneTabs {
    int32_t*  topo;      // [7]  NVLink TopologyScore of a uniform node for k = 2..8 (uniform_topo_build)
    uint32_t* grp;       // [6]  numa_pick lo/hi, numa_order, pcie_pick lo/hi, pcie_order
    uint32_t* mig;       // [3]  mig_free[0..9] as bytes
    int2*     memo;      // [7]  memoised winner per k: (TopologyScore or -1, join order)
    int4*     mat;       // [32] 16-byte units: int32 rows (unit 2m + half) or float64 rows (unit 4m + q)
    uint32_t* subset;    // [SUBSET_UNITS] 16-byte units of E1 subset keys (the pointer is to the lane's first unit)
    int       stride;
};

__device__ __forceinline__ uint32_t bfind_u32(uint32_t x) {
#ifdef __CUDA_ARCH__
    uint32_t r; asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x)); return r;
#else
    return x ? 31u - (uint32_t)__builtin_clz(x) : 0xFFFFFFFFu;
#endif
}

// PRMT as the hardware defines it (bit 3 of a selector nibble replicates the sign; __byte_perm masks the selector
// with 0x7777 first, one more instruction per step).  Selectors here are 0..7, or all ones on a dead walk,
// whose result is ANDed with an empty set anyway.
__device__ __forceinline__ uint32_t prmt_raw(uint32_t a, uint32_t b, uint32_t sel) {
#ifdef __CUDA_ARCH__
    uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel)); return d;
#else
    const uint64_t src = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        const uint32_t n = (sel >> (4 * i)) & 0xfu;
        uint32_t byte = (uint32_t)(src >> (8 * (n & 7u))) & 0xffu;
        if (n & 8u) byte = (byte & 0x80u) ? 0xffu : 0u;
        r |= byte << (8 * i);
    }
    return r;
#endif
}

// ---- uniform tier -----------------------------------------------------------------------------------
// All eight starts.  A start that is not available has an all-zero row, so its walk dies at once.  A walk with K
// members is K-1 picks; a dead set stays dead (x & 0), so one test before the last pick decides whether the walk
// completed.  All completed walks have the same total (uniform_topo_build), and the reference keeps the FIRST
// start that reaches the best total (strict >, scheduler.go:428-431): the starts are walked from GPU 7 down to
// GPU 0 and every completed walk overwrites the result, so the lowest-numbered completed start is what is left.
// ordN = join order in reversed positions, one NIBBLE per member (member p in bits [4p, 4p+4)).
template <int K>
__device__ __forceinline__ void walk_all(uint32_t rlo, uint32_t rhi, bool& found, uint32_t& ordN) {
    found = false; ordN = 0;
#pragma unroll
    for (int b0 = 0; b0 <= 7; ++b0) {
        uint32_t alive = prmt_raw(b0 < 4 ? rlo : rhi, 0u, 0x4440u | (uint32_t)(b0 & 3));      // byte b0 of the rows, zero-extended
        uint32_t ord = (uint32_t)b0;
#pragma unroll
        for (int t = 1; t < K - 1; ++t) {
            const uint32_t c = bfind_u32(alive);
            ord = c * (1u << (4 * t)) + ord;
            alive &= prmt_raw(rlo, rhi, c);
        }
        const bool ok = alive != 0u;
        const uint32_t c = bfind_u32(alive);
        ord = c * (1u << (4 * (K - 1))) + ord;
        ordN = ok ? ord : ordN;
        found = found || ok;
    }
}

__device__ __forceinline__ void walk_all_k(uint32_t rlo, uint32_t rhi, int k, bool& found, uint32_t& ordN) {
    switch (k) {      // k is the same for the whole warp
        case 2: walk_all<2>(rlo, rhi, found, ordN); break;
        case 3: walk_all<3>(rlo, rhi, found, ordN); break;
        case 4: walk_all<4>(rlo, rhi, found, ordN); break;
        case 5: walk_all<5>(rlo, rhi, found, ordN); break;
        case 6: walk_all<6>(rlo, rhi, found, ordN); break;
        case 7: walk_all<7>(rlo, rhi, found, ordN); break;
        case 8: walk_all<8>(rlo, rhi, found, ordN); break;
        default: found = false; ordN = 0; break;
    }
}

// KGWE_CFG_PRUNE_TIED_STARTS: only the walk from the lowest available start.  When it completes it is the
// reference's answer (all completed walks tie, the first start wins); when it dies the caller walks all starts.
__device__ __forceinline__ bool walk_first(uint32_t avail, uint32_t rlo, uint32_t rhi, int k, uint32_t& ordN) {
    const uint32_t b0 = 7u - (uint32_t)(__ffs((int)avail) - 1);
    uint32_t alive = prmt_raw(rlo, rhi, b0) & 0xffu;
    uint32_t ord = b0;
    int t = 1;
    for (; t < k && alive; ++t) {
        const uint32_t c = bfind_u32(alive);
        ord |= c << (4 * t);
        alive &= prmt_raw(rlo, rhi, c);
    }
    ordN = ord;
    return t == k;
}

// Nibble order in reversed positions -> SelectedGPUs as a set (GPU g at bit g) and as the ABI's packed list
// (3 bits per member, GPU indices).  k in 2..8.  The set: each nibble selects its byte 1 << (7 - b) out of an
// 8-byte table with one PRMT per four members; nibbles past k are set to 9, which PRMT reads as "sign of table
// byte 1" = 0.  The list: 7 - b = b ^ 7, then the nibbles are squeezed to 3-bit fields.
__device__ __forceinline__ void order_from_nibbles(uint32_t ordN, int k, uint32_t& sel, uint32_t& order) {
    const uint32_t used = k >= 8 ? 0xffffffffu : ((1u << (4 * k)) - 1u);
    const uint32_t f = (ordN & used) | (0x99999999u & ~used);
    uint32_t y = prmt_raw(0x10204080u, 0x01020408u, f) | prmt_raw(0x10204080u, 0x01020408u, f >> 16);
    y |= y >> 16;
    sel = (y | (y >> 8)) & 0xffu;
    uint32_t x = (ordN ^ 0x77777777u) & used;
    x = (x & 0x07070707u) | ((x >> 1) & 0x38383838u);       // two members per byte
    x = (x & 0x003f003fu) | ((x >> 2) & 0x0fc00fc0u);       // four per halfword
    order = (x & 0x00000fffu) | ((x >> 4) & 0x00fff000u);
}

// ---- int32 tier: the lane's matrix column ---------------------------------------------------------
__device__ __forceinline__ void lane_mat_build_int(const LaneTabs& tb, const DevNode* rec, uint64_t colact) {
    for (int m = 0; m < 8; ++m) {
        const uint32_t act = (uint32_t)(colact >> (8 * m)) & 0xffu;
        int v[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const double b = rec->bw_in[m][bw_slot(m, c)];
            v[c] = ((act >> c & 1) && c != m) ? (__double2int_rz(b) << FAST_SHIFT) : FAST_NEG;
        }
        tb.mat[(2 * m) * tb.stride] = make_int4(v[0], v[1], v[2], v[3]);
        tb.mat[(2 * m + 1) * tb.stride] = make_int4(v[4], v[5], v[6], v[7]);
    }
}
__device__ __forceinline__ void lane_mat_build_f64(const LaneTabs& tb, const DevNode* rec) {
    for (int m = 0; m < 8; ++m)
        for (int q = 0; q < 4; ++q) {
            const double2 v = reinterpret_cast<const double2*>(rec->bw_in[m])[q ^ (m >> 1)];      // (bw(2q -> m), bw(2q+1 -> m))
            int4 w;
            w.x = (int)(uint32_t)(unsigned long long)__double_as_longlong(v.x); w.y = (int)(uint32_t)((unsigned long long)__double_as_longlong(v.x) >> 32);
            w.z = (int)(uint32_t)(unsigned long long)__double_as_longlong(v.y); w.w = (int)(uint32_t)((unsigned long long)__double_as_longlong(v.y) >> 32);
            tb.mat[(4 * m + q) * tb.stride] = w;
        }
}

// All available starts, integer chains (see NodeFast / greedy_chain_fast_k): key = (total << 3) | (7 - start) of
// the best chain that reached K members with total > 0, else 0; pay = its join order.  Two starts are grown side by
// side (independent dependency chains: load -> add -> max tree -> next row), the second one masked when the number
// of available starts is odd.
template <int K>
__device__ __forceinline__ void chain_int_all(const int4* mat, int stride, uint32_t avail, int32_t& key, uint32_t& pay) {
    key = 0; pay = 0;
    int4 ia, ib;
    ia.x = (avail & 1u) ? 7 : FAST_NEG;   ia.y = (avail & 2u) ? 6 : FAST_NEG;   ia.z = (avail & 4u) ? 5 : FAST_NEG;   ia.w = (avail & 8u) ? 4 : FAST_NEG;
    ib.x = (avail & 16u) ? 3 : FAST_NEG;  ib.y = (avail & 32u) ? 2 : FAST_NEG;  ib.z = (avail & 64u) ? 1 : FAST_NEG;  ib.w = (avail & 128u) ? 0 : FAST_NEG;
#pragma unroll 1
    for (uint32_t rest = avail; rest;) {
        const int s0 = __ffs((int)rest) - 1; rest &= rest - 1u;
        const bool two = rest != 0u;
        const int s1 = two ? __ffs((int)rest) - 1 : s0; rest &= rest - 1u;
        int4 a0 = ia, b0 = ib, a1 = ia, b1 = ib;
        uint32_t sum0 = 0, sum1 = 0, ord0 = (uint32_t)s0, ord1 = (uint32_t)s1; int m0 = s0, m1 = s1;
        bool alive0 = true, alive1 = two;
#pragma unroll
        for (int step = 1; step < K; ++step) {
            const int4 r0 = mat[(2 * m0) * stride], r1 = mat[(2 * m0 + 1) * stride];
            const int4 q0 = mat[(2 * m1) * stride], q1 = mat[(2 * m1 + 1) * stride];
            a0.x += r0.x; a0.y += r0.y; a0.z += r0.z; a0.w += r0.w; b0.x += r1.x; b0.y += r1.y; b0.z += r1.z; b0.w += r1.w;
            a1.x += q0.x; a1.y += q0.y; a1.z += q0.z; a1.w += q0.w; b1.x += q1.x; b1.y += q1.y; b1.z += q1.z; b1.w += q1.w;
            const int32_t best0 = __vimax3_s32(__vimax3_s32(a0.x, a0.y, a0.z), __vimax3_s32(a0.w, b0.x, b0.y), max(b0.z, b0.w));
            const int32_t best1 = __vimax3_s32(__vimax3_s32(a1.x, a1.y, a1.z), __vimax3_s32(a1.w, b1.x, b1.y), max(b1.z, b1.w));
            alive0 &= best0 > 7; alive1 &= best1 > 7;
            m0 = ~best0 & 7; m1 = ~best1 & 7;
            sum0 += (uint32_t)best0; sum1 += (uint32_t)best1;
            ord0 |= (uint32_t)m0 << (3 * step); ord1 |= (uint32_t)m1 << (3 * step);
        }
        const int32_t k0 = alive0 ? (int32_t)(((sum0 >> (FAST_SHIFT - 3)) & ~7u) | (uint32_t)(7 - s0)) : 0;
        const int32_t k1 = alive1 ? (int32_t)(((sum1 >> (FAST_SHIFT - 3)) & ~7u) | (uint32_t)(7 - s1)) : 0;
        if (k0 > key) { key = k0; pay = ord0; }
        if (k1 > key) { key = k1; pay = ord1; }
    }
}
__device__ __forceinline__ void chain_int_all_k(const int4* mat, int stride, uint32_t avail, int k, int32_t& key, uint32_t& pay) {
    switch (k) {
        case 2: chain_int_all<2>(mat, stride, avail, key, pay); break;
        case 3: chain_int_all<3>(mat, stride, avail, key, pay); break;
        case 4: chain_int_all<4>(mat, stride, avail, key, pay); break;
        case 5: chain_int_all<5>(mat, stride, avail, key, pay); break;
        case 6: chain_int_all<6>(mat, stride, avail, key, pay); break;
        case 7: chain_int_all<7>(mat, stride, avail, key, pay); break;
        case 8: chain_int_all<8>(mat, stride, avail, key, pay); break;
        default: key = 0; pay = 0; break;
    }
}

// ---- float64 tier: greedy_chain (kgwe_device.cuh) on the lane's column, all available starts ----------
__device__ __forceinline__ double2 unit_as_double2(const int4& w) {
    double2 v;
    v.x = __longlong_as_double((long long)(((unsigned long long)(uint32_t)w.y << 32) | (uint32_t)w.x));
    v.y = __longlong_as_double((long long)(((unsigned long long)(uint32_t)w.w << 32) | (uint32_t)w.z));
    return v;
}
__device__ inline void chain_f64_all(const int4* mat, int stride, uint32_t avail, uint64_t colact, int k,
                                     bool& found, double& win_total, uint32_t& win_order) {
    found = false; win_total = 0.0; win_order = 0;
#pragma unroll 1
    for (uint32_t rest = avail; rest; rest &= rest - 1u) {
        const int s = __ffs((int)rest) - 1;
        double agg[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) agg[c] = 0.0;
        uint32_t members = 1u << s, conn = avail & ~members, order = (uint32_t)s;
        double total = 0.0; int m = s; bool ok = true;
        for (int step = 1; step < k; ++step) {
            conn &= (uint32_t)(colact >> (8 * m)) & 0xffu;
            double best = 0.0; int bc = -1;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double2 v = unit_as_double2(mat[(4 * m + q) * stride]);
                agg[2 * q] = __dadd_rn(agg[2 * q], v.x);
                agg[2 * q + 1] = __dadd_rn(agg[2 * q + 1], v.y);
                if ((conn >> (2 * q) & 1) && agg[2 * q] > best) { best = agg[2 * q]; bc = 2 * q; }
                if ((conn >> (2 * q + 1) & 1) && agg[2 * q + 1] > best) { best = agg[2 * q + 1]; bc = 2 * q + 1; }
            }
            if (bc < 0) { ok = false; break; }
            members |= 1u << bc; conn &= ~(1u << bc);
            order |= (uint32_t)bc << (3 * step);
            total = __dadd_rn(total, best);
            m = bc;
        }
        if (ok && total > win_total) { found = true; win_total = total; win_order = order; }      // strict >: the first start keeps ties
    }
}
// The same from the record itself (kernel variant without matrix columns; greedy_chain reads rec->bw_in).
__device__ inline void chain_f64_all_rec(const DevNode* rec, uint32_t avail, uint64_t colact, int k,
                                         bool& found, double& win_total, uint32_t& win_order) {
    found = false; win_total = 0.0; win_order = 0;
    NodeCtx nc; nc.avail = avail; nc.colact = colact;
#pragma unroll 1
    for (uint32_t rest = avail; rest; rest &= rest - 1u) {
        const int s = __ffs((int)rest) - 1;
        bool ok; double total; uint32_t order, members;
        greedy_chain(rec, nc, s, k, ok, total, order, members);
        if (ok && total > win_total) { found = true; win_total = total; win_order = order; }
    }
}

// ---- E1 exhaustive subsets ------------------------------------------------------------------------
// Subset keys of a node sit in (popcount, mask) order, every popcount class padded with zero keys to a multiple of four
// entries, four entries to a 16-byte unit, units laid out [unit][lane]: an evaluation for group size k reads the units of
// class k with conflict-free 128-bit loads.
constexpr int SUBSET_ENTRIES = 264;               // 4 + 8 + 28 + 56 + 72 + 56 + 28 + 8 + 4
constexpr int SUBSET_UNITS = SUBSET_ENTRIES / 4;
__host__ __device__ inline int subset_off(int k) {      // first entry of popcount class k (a multiple of 4)
    return k <= 0 ? 0 : k == 1 ? 4 : k == 2 ? 12 : k == 3 ? 40 : k == 4 ? 96 : k == 5 ? 168 : k == 6 ? 224 : k == 7 ? 252 : k == 8 ? 260 : 264;
}
// entry e of this lane's column, as a 32-bit word
__device__ __forceinline__ uint32_t& subset_entry(uint32_t* col, int S, int e) { return col[((e >> 2) * S) * 4 + (e & 3)]; }
// Subset keys of an integral node: key(mask) = total << 8 | 255 - mask when mask is an available clique on ACTIVE
// links (both directions of every pair) with total > 0, else 0; total = sum over pairs i < j of bw(j -> i).
// Built by subset extension: mask = rest + its highest GPU t adds sum_{i in rest} bw(t -> i).  slot_of[mask] is
// the entry of the mask.  First pass leaves (total << 1 | valid) in the entries, the second pass keys.
__device__ inline void lane_subset_build(const LaneTabs& tb, const LaneNode& ln, uint64_t colact, const uint16_t* slot_of) {
    const int S = tb.stride;
    for (int u = 0; u < SUBSET_UNITS; ++u) reinterpret_cast<uint4*>(tb.subset)[u * S] = make_uint4(0u, 0u, 0u, 0u);
    if (!ln.integral || !ln.valid) return;
    const int* mi = reinterpret_cast<const int*>(tb.mat);
    uint32_t sym[8];       // GPUs with an active link to t AND from t
    for (int t = 0; t < 8; ++t) {
        uint32_t out_t = 0;
        for (int i = 0; i < 8; ++i) if ((uint32_t)(colact >> (8 * i)) >> t & 1u) out_t |= 1u << i;
        sym[t] = (uint32_t)(colact >> (8 * t)) & 0xffu & out_t;
    }
    subset_entry(tb.subset, S, slot_of[0]) = 1u;                       // empty set: total 0, valid
    for (uint32_t mask = 1; mask < 256; ++mask) {
        const int t = 31 - __clz((int)mask);
        const uint32_t rest = mask & ~(1u << t);
        const uint32_t pr = subset_entry(tb.subset, S, slot_of[rest]);
        uint32_t ok = (pr & 1u) && !(rest & ~sym[t]) && (ln.avail >> t & 1u);
        uint32_t tot = pr >> 1;
        if (ok)
            for (uint32_t r = rest; r; r &= r - 1u) {
                const int i = __ffs((int)r) - 1;
                // unit 2i + (t >> 2), component t & 3: 64 * bw(t -> i) (active, so not FAST_NEG)
                tot += (uint32_t)(mi[((2 * i + (t >> 2)) * S) * 4 + (t & 3)] >> FAST_SHIFT);
            }
        subset_entry(tb.subset, S, slot_of[mask]) = ok ? ((tot << 1) | 1u) : 0u;
    }
    for (uint32_t mask = 0; mask < 256; ++mask) {
        uint32_t& e = subset_entry(tb.subset, S, slot_of[mask]);
        const uint32_t v = e;
        e = ((v & 1u) && (v >> 1) > 0u) ? (((v >> 1) << 8) | (255u - mask)) : 0u;
    }
}
// C(A, k) for 0 <= k <= A <= 8: Pascal's triangle, row A in one 64-bit word, 7 bits per entry (the largest is 70).
__device__ __forceinline__ uint32_t binom_small(int A, int k) {
    uint64_t row = 1, prev = 1;                               // row 0
    for (int a = 1; a <= A; ++a) { row = prev + (prev << 7); prev = row; }      // row a = row a-1 + (row a-1 shifted by one entry); no carries: entries < 128
    return (uint32_t)(row >> (7 * k)) & 0x7fu;
}

// ---- the evaluation ---------------------------------------------------------------------------------
constexpr uint32_t K1_MODE_FULL = 0;     // every available start GPU grows its group (the reference's search)
constexpr uint32_t K1_MODE_PRUNE = 1;    // KGWE_CFG_PRUNE_TIED_STARTS: uniform nodes stop at the first completed walk
constexpr uint32_t K1_MODE_MEMO = 2;     // KGWE_CFG_MEMOISE_GROUPS: winners per (node, k) found once per CTA

// The NVLink group search of one lane for group size k (2..8), not exhaustive.  topo = TopologyScore of the
// winning group, -1 when no start reaches k members; sel / order = its GPUs.  scored = groups actually grown.
template <bool GENERAL>
__device__ __forceinline__ void lane_group_search(const LaneNode& ln, const LaneTabs& tb, const DevNode* rec, int k, double norm,
                                                  uint32_t mode, int& topo, uint32_t& sel, uint32_t& order, uint32_t& scored) {
    topo = -1; sel = 0; order = 0; scored = 0;
    if (ln.tier == TIER_UNIFORM) {
        bool found = false, all_starts = true; uint32_t ordN = 0;
        if (mode == K1_MODE_PRUNE) { found = walk_first(ln.avail, ln.rlo, ln.rhi, k, ordN); scored = 1; all_starts = !found; }
        if (all_starts) { walk_all_k(ln.rlo, ln.rhi, k, found, ordN); scored += (uint32_t)ln.A; }      // (one inlined copy of the walks in the pod loop)
        order_from_nibbles(ordN, k, sel, order);
        if (found) topo = tb.topo[(k - 2) * tb.stride];
    } else if (GENERAL && ln.tier == TIER_INT) {
        int32_t key; uint32_t pay;
        chain_int_all_k(tb.mat, tb.stride, ln.avail, k, key, pay); scored = (uint32_t)ln.A;
        if (key > 0) { topo = nvlink_score((double)(key >> 3), k, norm); order = pay; sel = mask_of_order(pay, k); }
    } else if (ln.tier != TIER_NONE) {
        bool found; double total; uint32_t ord;
        const uint64_t colact = ld_u64(rec->act_in);
        if (GENERAL && ln.tier == TIER_F64) chain_f64_all(tb.mat, tb.stride, ln.avail, colact, k, found, total, ord);
        else chain_f64_all_rec(rec, ln.avail, colact, k, found, total, ord);
        scored = (uint32_t)ln.A;
        if (found) { topo = nvlink_score(total, k, norm); order = ord; sel = mask_of_order(ord, k); }
    }
}

__device__ __forceinline__ bool lane_arch_match(const LaneNode& ln, uint32_t arch) {      // arch != 0; unavailable GPUs hold 0
    const uint32_t rep = arch * 0x01010101u;
    const uint32_t x = ln.arch_lo ^ rep, y = ln.arch_hi ^ rep;
    return ((((x - 0x01010101u) & ~x) | ((y - 0x01010101u) & ~y)) & 0x80808080u) != 0u;      // some byte is zero
}

// One evaluation's result: the 16-byte NodeScore as four words (all zero when the node is not eligible), the
// arg-max contribution (TotalScore when the evaluation is kept, else 0), and the two candidate counts.
struct LaneResult {
    uint32_t w0, w1, w2, w3;       // total | topology, resource << 16 | balance, gpu_mask << 16, flags << 24 | order
    uint32_t kept_total;           // TotalScore if KGWE_F_KEPT else 0 (kept implies TotalScore > 0)
    uint32_t candidates, scored;   // reference-semantic count / subsets whose score this call computed
};

// isNodeEligible + scoreNode (scheduler.go:206-300) for (this lane's node, pod).  `rec` is the node's record in
// global memory (read only on rare paths: minimum-bandwidth filter, float64 fall-backs).
template <bool GENERAL, bool TABLES>
__device__ __forceinline__ void lane_evaluate(const LaneNode& ln, const LaneTabs& tb, const DevNode* rec, const PodLite& pd,
                                              double min_bw, bool hinted, const Params& P, uint32_t mode, LaneResult& r) {
    r.w0 = r.w1 = r.w2 = r.w3 = 0u; r.kept_total = 0u; r.candidates = 1u; r.scored = 1u;
    const uint32_t pk = pd.pk;
    const int k = (int)(pk & 0xffu);
    bool eligible = ln.valid && ln.A >= k;                                            // scheduler.go:213-216
    if (pk & PL_ARCH) eligible = eligible && lane_arch_match(ln, (pk >> 16) & 0xffu);  // :219-230
    if (pk & PL_MIG) {                                                                // :233-238
        const uint32_t prof = pk >> 24;
        int have = 0;
        if (prof >= 1 && prof <= KGWE_MIG_PROFILES && ln.valid) {
            const uint32_t idx = prof - 1;
            have = (int)((tb.mig[(idx >> 2) * tb.stride] >> (8 * (idx & 3))) & 0xffu);
        }
        eligible = eligible && have >= pd.mig_count;
    }
    if (!eligible) return;
    int topo; uint32_t sel, order, fb = 0;       // fb: KGWE_F_FALLBACK or 0
    int nsel = k;
    // first k available GPUs in index order (k = 1 for k <= 1): the fall-back selection of every arm; kk <= A <= 8 (or A = 0: both are 0)
    const int kk = k < 1 ? 1 : k;
#define KGWE_SEL_FIRST ((uint32_t)(ln.avail_prefix >> (8 * (kk - 1))) & 0xffu)
#define KGWE_ORD_FIRST (ln.avail_order & ((1u << (3 * kk)) - 1u))
    if (k <= 1) {                                                                     // :309-315 (k = 0 lands here too)
        const bool any = ln.A > 0;
        topo = any ? 100 : 0; sel = KGWE_SEL_FIRST; order = KGWE_ORD_FIRST; nsel = any ? 1 : 0;
    } else if (pk & PL_NVLINK) {                                                      // scoreNVLinkTopology, :336-373
        int found_topo = -1;
        if (pk & PL_EXH) {
            r.candidates = binom_small(ln.A, k); r.scored = r.candidates;
            uint32_t bm = 0; double best = 0.0;
            if (TABLES && ln.integral) {
                uint32_t key = 0;
                const uint4* units = reinterpret_cast<const uint4*>(tb.subset);
#pragma unroll 2
                for (int u = subset_off(k) >> 2; u < (subset_off(k + 1) >> 2); ++u) {
                    const uint4 v = units[u * tb.stride];
                    key = max(max(key, max(v.x, v.y)), max(v.z, v.w));
                }
                if (key) { best = (double)(key >> 8); bm = 255u - (key & 255u); }
            } else {
                NodeCtx nc; nc.avail = ln.avail; nc.colact = ld_u64(rec->act_in);
                for (int j = 0; j < 8; ++j) {
                    double b; uint32_t m;
                    exhaustive_lane(rec, nc, j, k, b, m);
                    if (b > best || (b == best && b > 0.0 && m < bm)) { best = b; bm = m; }
                }
            }
            sel = 0; order = 0;
            if (best > 0.0) { found_topo = nvlink_score(best, k, P.norm); take_first_k(bm, k, sel, order); }
        } else {
            r.candidates = (uint32_t)ln.A;
            if (mode == K1_MODE_MEMO) {
                const int2 w = tb.memo[(k - 2) * tb.stride];
                found_topo = w.x; order = (uint32_t)w.y & 0xffffffu; sel = (uint32_t)w.y >> 24; r.scored = 0;
            } else lane_group_search<GENERAL>(ln, tb, rec, k, P.norm, mode, found_topo, sel, order, r.scored);
        }
        const bool found = found_topo >= 0;                                            // else no fully connected group, :358-365
        topo = found_topo;
        if (!found) { topo = 30; fb = KGWE_F_FALLBACK; sel = KGWE_SEL_FIRST; order = KGWE_ORD_FIRST; }
    } else if (pk & (PL_NUMA | PL_PCIE)) {                                            // :438-513
        const bool numa = pk & PL_NUMA;
        const int base = numa ? 0 : 3;
        const uint64_t pick = (uint64_t)tb.grp[base * tb.stride] | ((uint64_t)tb.grp[(base + 1) * tb.stride] << 32);
        const uint32_t gorder = tb.grp[(base + 2) * tb.stride];
        const uint32_t grp = (uint32_t)((pick >> (8 * (k - 1))) & 0xff);              // 2 <= k <= A <= 8 here
        if (grp) { topo = numa ? 90 : 80; take_first_k(grp, k, sel, order); }
        else {
            topo = numa ? 50 : 40; fb = KGWE_F_FALLBACK;
            order = gorder & ((1u << (3 * k)) - 1u);
            sel = mask_of_order(order, k);
        }
    } else {                                                                          // :325-331
        topo = 50; sel = KGWE_SEL_FIRST; order = KGWE_ORD_FIRST;
    }
#undef KGWE_SEL_FIRST
#undef KGWE_ORD_FIRST
    int res = 0;                                                                      // calculateResourceScore, :516-553
    if (ln.A > 0) {
        res = 50 + (ln.sum_free >= pd.required * 2 ? 25 : (ln.sum_free >= pd.required ? 10 : 0)) + ln.util_bonus;
        res = res > 100 ? 100 : res;
    }
    int total = (topo * P.wt + res * P.wr + ln.balance * P.wb) / 100;                 // :282-286
    uint32_t flags = KGWE_F_ELIGIBLE | fb;
    if (hinted) { total += P.bonus; flags |= KGWE_F_HINT; }
    bool kept = total > 0;                                                            // :196
    if ((pk & PL_MINBW) && kept && est_bandwidth(rec, order, nsel) < min_bw) { kept = false; flags |= KGWE_F_MINBW; }      // E3
    if (kept) flags |= KGWE_F_KEPT;
    r.w0 = (uint32_t)total;
    r.w1 = ((uint32_t)topo & 0xffffu) | ((uint32_t)res << 16);
    r.w2 = (uint32_t)ln.balance | (sel << 16) | (flags << 24);
    r.w3 = order | ((uint32_t)nsel << 24);
    r.kept_total = kept ? (uint32_t)total : 0u;
}

// What the pods of a chunk need from the per-node tables (a bit set, OR-ed over the chunk): bit k (2..8) = an NVLink pod with
// GPUCount k (NVLink score of the uniform node / memoised winner for that k), NEED_GROUPS = a SameNUMA / SamePCIeSwitch pod,
// NEED_MIG = a pod with MIG requirements.  A single-pod call builds one table entry instead of seven.
constexpr uint32_t NEED_GROUPS = 1u << 9, NEED_MIG = 1u << 10, NEED_ALL = 0xffffffffu;
__device__ __forceinline__ uint32_t pod_needs(const PodLite& pl) {
    uint32_t n = 0;
    const uint32_t k = pl.pk & 0xffu;
    if ((pl.pk & PL_NVLINK) && k >= 2u && k <= 8u) n |= 1u << k;
    if (pl.pk & (PL_NUMA | PL_PCIE)) n |= NEED_GROUPS;
    if (pl.pk & PL_MIG) n |= NEED_MIG;
    return n;
}

// Memoised winners of this lane's node for the needed k in 2..8 (K1_MODE_MEMO), by the full search: (TopologyScore or -1,
// join order | GPU set << 24).
template <bool GENERAL>
__device__ __noinline__ void lane_memo_build(const LaneNode& ln, const LaneTabs& tb, const DevNode* rec, double norm, uint32_t needs) {
    for (int k = 2; k <= 8; ++k) {
        if (!(needs >> k & 1u)) continue;
        int topo = -1; uint32_t sel = 0, order = 0, sc = 0;
        if (ln.valid && ln.A >= k) lane_group_search<GENERAL>(ln, tb, rec, k, norm, K1_MODE_FULL, topo, sel, order, sc);
        tb.memo[(k - 2) * tb.stride] = make_int2(topo, (int)(order | (sel << 24)));
    }
}

// Everything a lane keeps about its node beyond LaneNode: tables in the shared-memory columns.
template <bool GENERAL, bool TABLES>
__device__ inline void lane_tables_build(const LaneNode& ln, const NodeCtx& c, const LaneTabs& tb, const DevNode* rec, double norm,
                                         uint32_t mode, const uint16_t* slot_of, uint32_t needs = NEED_ALL) {
    const int S = tb.stride;
    if (ln.tier == TIER_UNIFORM) {                       // uniform_topo_build, into the lane's column
        const double u = node_uniform_bw(rec, node_uniform_code(rec));
        double Sm = 0.0, T = 0.0;
        for (int k = 2; k <= 8; ++k) {
            Sm = __dadd_rn(Sm, u); T = __dadd_rn(T, Sm);
            if (needs >> k & 1u) tb.topo[(k - 2) * S] = nvlink_score(T, k, norm);
        }
    }
    if (needs & NEED_GROUPS) {
        tb.grp[0] = (uint32_t)c.numa_pick; tb.grp[S] = (uint32_t)(c.numa_pick >> 32); tb.grp[2 * S] = c.numa_order;
        tb.grp[3 * S] = (uint32_t)c.pcie_pick; tb.grp[4 * S] = (uint32_t)(c.pcie_pick >> 32); tb.grp[5 * S] = c.pcie_order;
    }
    if (ln.valid && (needs & NEED_MIG)) {
        uint32_t w[3] = {0, 0, 0};
        for (int i = 0; i < KGWE_MIG_PROFILES; ++i) w[i >> 2] |= (uint32_t)rec->mig_free[i] << (8 * (i & 3));
        tb.mig[0] = w[0]; tb.mig[S] = w[1]; tb.mig[2 * S] = w[2];
    }
    if (GENERAL && ln.valid) {
        if (ln.integral) lane_mat_build_int(tb, rec, c.colact);
        else if (ln.tier == TIER_F64) lane_mat_build_f64(tb, rec);
    }
    if (TABLES) lane_subset_build(tb, ln, c.colact, slot_of);
    if (mode == K1_MODE_MEMO) lane_memo_build<GENERAL>(ln, tb, rec, norm, needs);
}

}  // namespace kgwe
