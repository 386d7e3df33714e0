// kgwe_device.cuh — device-side record layout and the per-(pod,node) evaluation shared by
// every kernel of the engine (stateless batch, stateful queue, finalisation).
//
// Execution shape: one (pod,node) evaluation is done by a group of 8 lanes, 4 groups per
// warp.  Lane j of a group owns GPU j of the node: it is the START GPU of one greedy
// NVLink chain (reference src/scheduler/scheduler.go:389), so the 8 chains of
// findBestNVLinkGroup run side by side and a 3-step segmented shuffle picks the winner
// with the reference's strict-greater / first-start rule.  Everything else in an
// evaluation is group-uniform integer work.
//
// Arithmetic: FP64 with explicit round-to-nearest intrinsics (__dadd_rn, __dmul_rn,
// __ddiv_rn) and truncating conversion, compiled with --fmad=false, because Go on amd64
// never contracts a*b+c and the integer scores must match bit for bit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/kgwe_engine.h"

namespace kgwe {

// Device image of one node: same 768 bytes as kgwe_node_image, with the NVLink matrix
// stored the way the greedy reads it.  bw_in[m] is the 64-byte row "bandwidth of GPU c's
// link to GPU m, for c = 0..7" (what a candidate c adds when m has joined the group);
// inside the row the 16-byte chunk holding c = 2q, 2q+1 sits at chunk position
// q ^ (m >> 1), so the 8 lanes of a group, each reading a different row m at the same
// logical chunk q, hit 8 distinct 4-bank groups of shared memory (conflict-free LDS.128).
struct __align__(16) DevNode {
    double   bw_in[8][8];
    double   pcie_bw[8];
    uint64_t free_bytes[8];
    double   util[8];
    uint8_t  act_in[8];      // act_in[m] bit c: GPU c has an ACTIVE link entry to GPU m
    uint8_t  present[8];     // present[i] bit j: GPU i has a link entry for GPU j (active or not)
    int8_t   numa[8];
    uint8_t  pcie_switch[8];
    uint8_t  arch[8];
    uint8_t  mig_free[KGWE_MIG_PROFILES];
    uint8_t  n_gpus, healthy_mask, allocated_mask, flags;
    uint8_t  mig_enabled_mask, mig_has_free_mask;
    uint32_t n_allocs;
    uint8_t  node_flags;
    uint8_t  pad1[3];
};
static_assert(sizeof(DevNode) == 768, "device record must stay 768 bytes");
static_assert(sizeof(kgwe_node_image) == 768, "ABI record must stay 768 bytes");
static_assert(sizeof(kgwe_pod_req) == 32 && sizeof(kgwe_node_score) == 16 && sizeof(kgwe_placement) == 32, "ABI sizes");

__host__ __device__ inline int bw_slot(int m, int c) { return (((c >> 1) ^ (m >> 1)) << 1) | (c & 1); }

// DevNode.flags
constexpr uint8_t NODE_PRESENT = 0x01;
constexpr uint8_t NODE_INTEGRAL = 0x02;   // every link bandwidth is a whole number of GB/s (set by the host packer)

// Integer form of one node's greedy inputs, built in shared memory by the scoring kernel for
// nodes whose bandwidths are whole numbers (<= KGWE_MAX_BW_GBPS < 2^17).  Sums of such doubles
// are exact in FP64, so int32 arithmetic yields the very same totals and the same comparisons
// as the reference's float64 code.  Values are pre-scaled by 64 and the candidate's tie-break
// code (7 - c) is folded into the low bits through init[], so `max` over the eight running
// sums IS the reference's "largest aggregate, first candidate on ties" rule, and the sum of the
// per-step maxima is 64 * total + (at most 7 * 7 of tie codes), i.e. total = sum >> 6:
//   row[m][c]  = 64 * bw(c -> m) when that link is active, FAST_NEG otherwise (and on the diagonal,
//                which retires a GPU from the candidates the step after it joins);
//   init[c]    = 7 - c for an available GPU, FAST_NEG otherwise.
// A running sum that ever met FAST_NEG stays negative (at most 8 of them: no int32 overflow;
// positive parts stay below 7 * 64 * 1e5 < 2^26).  Rows are 32 bytes at a 48-byte stride, which
// puts the 8 rows' 16-byte chunks on 8 distinct bank groups: 8 lanes on 8 rows never conflict.
constexpr int32_t FAST_NEG = -(1 << 27);
constexpr int FAST_SHIFT = 6;
struct __align__(16) NodeFast { int32_t row[8][12]; int32_t init[8]; };

struct Params {          // kgwe_config, device copy
    int32_t wt, wr, wb, bonus;
    double  norm;
    uint32_t node_base;
};

// Pod-independent facts about one node, recomputed only when its ledger fields change.
struct NodeCtx {
    uint32_t avail;        // healthy & ~allocated & (n_gpus bits)        scheduler.go:581-603
    int      A;
    uint64_t colact;       // byte m = act_in[m]
    uint64_t sum_free;     // sum of FreeBytes over available GPUs          scheduler.go:528-531
    int      util_bonus;   // 25 / 10 / 0 from the ordered mean of GPUPercent scheduler.go:540-550
    int      balance;      // calculateBalanceScore                          scheduler.go:556-578
    uint64_t arch8;        // arch[0..7]
    uint64_t numa_pick, pcie_pick;    // byte k-1: first group (ascending key) with >= k members, else 0
    uint32_t numa_order, pcie_order;  // available GPUs in (key asc, index asc) order, 3 bits each
    uint64_t avail_prefix;            // byte k-1: the first k available GPUs (index order) as a mask
    uint32_t avail_order;             // available GPUs in index order, 3 bits each
    bool     valid, integral, tables_valid;
};

struct EvalOut {
    int32_t  total, topo, res, bal;
    uint32_t mask, order, flags;   // order: packed list + count in bits 24..27
    uint32_t candidates;
};

__device__ __forceinline__ uint64_t ld_u64(const void* p) { return *reinterpret_cast<const uint64_t*>(p); }

// Groups of available GPUs by an 8-bit key, visited in ascending key (canonical order of
// scoreNUMATopology / scorePCIeTopology, scheduler.go:438-513).
__device__ inline void group_tables(uint64_t keys8, bool is_signed, uint32_t avail, uint64_t& pick, uint32_t& order) {
    pick = 0; order = 0;
    uint32_t remaining = avail; int pos = 0;
    while (remaining) {
        int minkey = 1 << 30;
        for (int g = 0; g < 8; ++g)
            if (remaining >> g & 1) {
                int kb = (int)((keys8 >> (8 * g)) & 0xff);
                int key = is_signed ? (int)(int8_t)kb : kb;
                minkey = key < minkey ? key : minkey;
            }
        uint32_t grp = 0;
        for (int g = 0; g < 8; ++g)
            if (remaining >> g & 1) {
                int kb = (int)((keys8 >> (8 * g)) & 0xff);
                int key = is_signed ? (int)(int8_t)kb : kb;
                if (key == minkey) { grp |= 1u << g; order |= (uint32_t)g << (3 * pos); ++pos; }
            }
        int sz = __popc(grp);
        for (int k = 1; k <= sz; ++k)
            if (((pick >> (8 * (k - 1))) & 0xff) == 0) pick |= (uint64_t)grp << (8 * (k - 1));
        remaining &= ~grp;
    }
}

__device__ inline void node_ctx_group_tables(NodeCtx& c, const DevNode* rec) {
    if (c.valid) {
        group_tables(ld_u64(rec->numa), true, c.avail, c.numa_pick, c.numa_order);
        group_tables(ld_u64(rec->pcie_switch), false, c.avail, c.pcie_pick, c.pcie_order);
    } else { c.numa_pick = c.pcie_pick = 0; c.numa_order = c.pcie_order = 0; }
    c.tables_valid = true;
}

// Everything of NodeCtx except the NUMA / PCIe group tables (tables_valid = false).
__device__ inline void node_ctx_core(NodeCtx& c, const DevNode* rec, bool in_range) {
    c.valid = in_range && (rec->flags & NODE_PRESENT);
    c.integral = c.valid && (rec->flags & NODE_INTEGRAL);
    uint32_t ng = c.valid ? rec->n_gpus : 0;
    ng = ng > 8 ? 8 : ng;
    uint32_t avail = c.valid ? ((uint32_t)rec->healthy_mask & ~(uint32_t)rec->allocated_mask & ((1u << ng) - 1u)) : 0;
    c.avail = avail; c.A = __popc(avail);
    c.colact = c.valid ? ld_u64(rec->act_in) : 0;
    c.arch8 = c.valid ? ld_u64(rec->arch) : 0;
    uint64_t sum = 0; double usum = 0.0;
    for (int g = 0; g < 8; ++g)
        if (avail >> g & 1) { sum += rec->free_bytes[g]; usum = __dadd_rn(usum, rec->util[g]); }
    c.sum_free = sum;
    c.util_bonus = 0;
    if (c.A > 0) {
        double avg = __ddiv_rn(usum, (double)c.A);
        c.util_bonus = avg < 30.0 ? 25 : (avg < 60.0 ? 10 : 0);
    }
    c.balance = 0;
    if (ng > 0) {
        double usage = __ddiv_rn((double)rec->n_allocs, (double)ng);
        int s = __double2int_rz(__dmul_rn(100.0, __dsub_rn(1.0, usage)));
        c.balance = s > 0 ? s : 0;
    }
    c.avail_prefix = 0; c.avail_order = 0;
    {
        uint32_t m = avail, acc = 0; int i = 0;
        while (m) {
            const int b = __ffs(m) - 1;
            acc |= 1u << b; c.avail_order |= (uint32_t)b << (3 * i); c.avail_prefix |= (uint64_t)acc << (8 * i);
            m &= m - 1; ++i;
        }
    }
    c.numa_pick = c.pcie_pick = 0; c.numa_order = c.pcie_order = 0; c.tables_valid = false;
}

__device__ inline void node_ctx_init(NodeCtx& c, const DevNode* rec, bool in_range) {
    node_ctx_core(c, rec, in_range);
    node_ctx_group_tables(c, rec);
}

// First k set bits of `m` in index order -> (mask, packed list).
__device__ __forceinline__ void take_first_k(uint32_t m, int k, uint32_t& sel, uint32_t& order) {
    sel = 0; order = 0;
    for (int i = 0; i < k && m; ++i) {
        int b = __ffs(m) - 1;
        sel |= 1u << b; order |= (uint32_t)b << (3 * i);
        m &= m - 1;
    }
}
// GPU mask of a packed list (k fields of 3 bits), without a loop: lanes of a warp hold lists of different
// lengths, so a loop runs as long as the longest one.  The unused fields are filled with field 0 (a member
// anyway), the eight fields are spread to nibbles, and each nibble then selects its own byte 1 << v out of
// the table 0x8040201008040201 with one PRMT per four fields; OR-ing the eight bytes gives the mask.
__device__ __forceinline__ uint32_t mask_of_order(uint32_t order, int k) {
    if (k <= 0) return 0u;
    const uint32_t used = k >= 8 ? 0x00ffffffu : ((1u << (3 * k)) - 1u);
    uint32_t x = (order & used) | (((order & 7u) * 0x249249u) & ~used & 0x00ffffffu);
    x = (x & 0x00000fffu) | ((x & 0x00fff000u) << 4);       // 4 fields per halfword
    x = (x & 0x003f003fu) | ((x & 0x0fc00fc0u) << 2);       // 2 fields per byte
    x = (x & 0x07070707u) | ((x & 0x38383838u) << 1);       // 1 field per nibble
    uint32_t y = __byte_perm(0x08040201u, 0x80402010u, x) | __byte_perm(0x08040201u, 0x80402010u, x >> 16);
    y |= y >> 16;
    return (y | (y >> 8)) & 0xffu;
}

// First min(k, A) available GPUs in index order, from the per-node tables.
__device__ __forceinline__ void first_k_avail(const NodeCtx& nc, int k, uint32_t& sel, uint32_t& order, int& n) {
    n = k < nc.A ? k : nc.A;
    if (n <= 0) { sel = 0; order = 0; n = 0; return; }
    sel = (uint32_t)(nc.avail_prefix >> (8 * (n - 1))) & 0xffu;
    order = nc.avail_order & (n >= 8 ? 0xffffffu : ((1u << (3 * n)) - 1u));
}

// Uniform nodes: every ACTIVE link of the node carries the same bandwidth u > 0 - any NVSwitch node, any single-generation
// mesh, whole-number or not (the host packer records where one such link sits in DevNode.pad1[0]: 0x80 | m << 3 | c,
// 0x40 = "no active link at all", 0 = not uniform).  Every candidate that is linked to all members then has the SAME
// aggregate (the same additions of u in the same order, so this holds bit for bit in float64 too), and "largest
// aggregate, first candidate on ties" is simply "lowest-numbered candidate still linked to all members".  A chain becomes
// a walk over bit sets held in registers:
//   alive = avail & act_in[s];   step: c = lowest bit of alive;  alive &= act_in[c]   (act_in[c] never contains c)
// and its total depends on (u, K) only: T[K] = S_1 + ... + S_(K-1) with S_t = S_(t-1) + u, added in that order in float64
// as the reference does (uniform_totals_build; exact k(k-1)/2 * u for whole numbers).  key = 8 | (7 - s) marks a walk
// that reached K members (the totals are equal, so the key only ranks the start lanes); pay = join order.
__device__ __forceinline__ int node_uniform_code(const DevNode* rec) { return rec->pad1[0]; }
__device__ __forceinline__ double node_uniform_bw(const DevNode* rec, int code) {
    return (code & 0x80) ? rec->bw_in[(code >> 3) & 7][bw_slot((code >> 3) & 7, code & 7)] : 0.0;
}

template <int K>
__device__ __forceinline__ void greedy_walk_uniform_k(uint32_t avail, uint64_t colact, int s, int32_t& key, uint32_t& pay) {
    const uint32_t lo = (uint32_t)colact, hi = (uint32_t)(colact >> 32);
    uint32_t alive = avail & __byte_perm(lo, hi, (uint32_t)s);       // byte s of colact = act_in[s]; avail <= 0xff masks the rest
    uint32_t order = (uint32_t)s;
    bool ok = true;
#pragma unroll
    for (int step = 1; step < K; ++step) {
        ok &= alive != 0u;
        const uint32_t c = (uint32_t)(__ffs((int)alive) - 1);        // 0..7; all ones once the walk is dead (alive == 0 stays 0, key dropped)
        order += c << (3 * step);
        alive &= __byte_perm(lo, hi, c);
    }
    key = ok ? (8 | (7 - s)) : 0;
    pay = order;
}

__device__ __forceinline__ void greedy_walk_uniform(uint32_t avail, uint64_t colact, int s, int k, int32_t& key, uint32_t& pay) {
    if (k < 5) {
        if (k < 3) { if (k == 2) greedy_walk_uniform_k<2>(avail, colact, s, key, pay); else { key = 0; pay = 0; } }
        else if (k == 3) greedy_walk_uniform_k<3>(avail, colact, s, key, pay);
        else greedy_walk_uniform_k<4>(avail, colact, s, key, pay);
    } else if (k < 7) {
        if (k == 5) greedy_walk_uniform_k<5>(avail, colact, s, key, pay); else greedy_walk_uniform_k<6>(avail, colact, s, key, pay);
    } else if (k == 7) greedy_walk_uniform_k<7>(avail, colact, s, key, pay);
    else if (k == 8) greedy_walk_uniform_k<8>(avail, colact, s, key, pay);
    else { key = 0; pay = 0; }
}

// One lane, one evaluation: the walk from the LOWEST available start.  All walks that reach k members have the same total
// and the reference keeps the first start that reaches the best total (strict >, scheduler.go:428-431), so when this walk
// completes it is the reference's answer and the other seven starts cannot change it.  When it dies (a link down, a
// sparse mesh) the caller falls back to the 8-lane search over all starts.  2 <= k <= 8, avail != 0.
__device__ __forceinline__ bool first_uniform_walk(uint32_t avail, uint64_t colact, int k, uint32_t& order_out) {
    const uint32_t lo = (uint32_t)colact, hi = (uint32_t)(colact >> 32);
    const uint32_t s = (uint32_t)(__ffs((int)avail) - 1);
    uint32_t alive = avail & __byte_perm(lo, hi, s);
    uint32_t order = s;
    int step = 1;
    for (; step < k && alive; ++step) {
        const uint32_t c = (uint32_t)(__ffs((int)alive) - 1);
        order += c << (3 * step);
        alive &= __byte_perm(lo, hi, c);
    }
    order_out = order;
    return step == k;
}

// The NVLink TopologyScore of a uniform node depends on k only: the walk's total is T[k] = S_1 + ... + S_(k-1) with
// S_t = S_(t-1) + u, added in that order in float64 as the reference does (exact k(k-1)/2 * u for whole numbers), and the
// score is nvlink_score(T[k]).  Built once per node per CTA, kept as 7 ints in the spare 16 bytes at the end of NodeFast
// rows 0 and 1, so an evaluation on a uniform node does no floating-point arithmetic at all.
__device__ __forceinline__ int* uniform_topo_slot(NodeFast* nf, int k) {             // 2 <= k <= 8
    return &nf->row[(k - 2) >> 2][8 + ((k - 2) & 3)];
}
__device__ __forceinline__ int nvlink_score(double total, int k, double norm);
__device__ __forceinline__ void uniform_topo_build(NodeFast* nf, double u, double norm) {
    double S = 0.0, T = 0.0;
    for (int k = 2; k <= 8; ++k) {
        S = __dadd_rn(S, u);            // the aggregate of the candidate chosen at step k-1: u added k-1 times from 0.0
        T = __dadd_rn(T, S);            // findBestNVLinkGroup's running total
        *uniform_topo_slot(nf, k) = nvlink_score(T, k, norm);
    }
}

// Lane j's build step of NodeFast: row j (what each candidate gains when GPU j joins).
__device__ __forceinline__ void node_fast_build(NodeFast* nf, const DevNode* rec, const NodeCtx& nc, int j) {
    const uint32_t act = (uint32_t)(nc.colact >> (8 * j)) & 0xffu;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const double v = rec->bw_in[j][bw_slot(j, c)];
        nf->row[j][c] = ((act >> c & 1) && c != j) ? (__double2int_rz(v) << FAST_SHIFT) : FAST_NEG;
    }
    nf->init[j] = (nc.avail >> j & 1) ? (7 - j) : FAST_NEG;
}

// Integer twin of greedy_chain (see NodeFast), fully unrolled for a group size K.  Returns
// key = (total << 3) | (7 - s) when the chain from start s reaches size K with total > 0, else 0;
// pay = join order (3 bits per position).
template <int K>
__device__ __forceinline__ void greedy_chain_fast_k(const NodeFast* nf, int s, int32_t& key, uint32_t& pay) {
    key = 0; pay = 0;
    const int4* ip = reinterpret_cast<const int4*>(nf->init);
    int4 a = ip[0], b = ip[1];                         // (the caller has checked that start s is available)
    uint32_t sum = 0; uint32_t order = (uint32_t)s; int m = s;      // unsigned: a dead chain may wrap, and must not be UB
    bool alive = true;
#pragma unroll
    for (int step = 1; step < K; ++step) {
        const int4* rp = reinterpret_cast<const int4*>(nf->row[m]);
        const int4 r0 = rp[0], r1 = rp[1];
        a.x += r0.x; a.y += r0.y; a.z += r0.z; a.w += r0.w;
        b.x += r1.x; b.y += r1.y; b.z += r1.z; b.w += r1.w;
        const int32_t best = __vimax3_s32(__vimax3_s32(a.x, a.y, a.z), __vimax3_s32(a.w, b.x, b.y), max(b.z, b.w));
        // best <= 7: nobody fully connected with a positive aggregate - the chain is dead.  No branch: dead chains are
        // rare, and a straight-line body lets the steps be scheduled as one block.  (A dead chain keeps stepping on
        // whatever m falls out; the running sums stay inside int32: at most 8 terms of FAST_NEG each.)
        alive &= best > 7;
        m = ~best & 7;                                  // 7 - (best & 7)
        sum += (uint32_t)best;
        order |= (uint32_t)m << (3 * step);
    }
    // An alive chain has total > 0 (every step's maximum was above 7, i.e. at least 1 GB/s), and
    // (sum >> FAST_SHIFT) << 3 is (sum >> 3) with the low three bits cleared.  pay is only read when key > 0.
    key = alive ? (int32_t)(((sum >> (FAST_SHIFT - 3)) & ~7u) | (uint32_t)(7 - s)) : 0;
    pay = order;
}

// The chain from start s for group size k is the first k-1 steps of the chain for size 8, so one
// run yields the outcome for every size: sums[t] = running sum of maxima after step t+1 (valid for
// t < steps), order = join order of the longest chain.  Used where a node's winners for all sizes
// are rebuilt at once (memo tables).
__device__ __forceinline__ void greedy_chain_fast_all(const NodeFast* nf, int s, int32_t (&sums)[7], int& steps, uint32_t& order) {
    steps = 0; order = (uint32_t)s;
    const int4* ip = reinterpret_cast<const int4*>(nf->init);
    int4 a = ip[0], b = ip[1];
    if (nf->init[s] < 0) { steps = -1; return; }
    int32_t sum = 0; int m = s;
#pragma unroll
    for (int step = 1; step < 8; ++step) {
        const int4* rp = reinterpret_cast<const int4*>(nf->row[m]);
        const int4 r0 = rp[0], r1 = rp[1];
        a.x += r0.x; a.y += r0.y; a.z += r0.z; a.w += r0.w;
        b.x += r1.x; b.y += r1.y; b.z += r1.z; b.w += r1.w;
        const int32_t best = __vimax3_s32(__vimax3_s32(a.x, a.y, a.z), __vimax3_s32(a.w, b.x, b.y), max(b.z, b.w));
        if (best <= 7) return;
        m = ~best & 7;
        sum += best;
        sums[step - 1] = sum; steps = step;
        order |= (uint32_t)m << (3 * step);
    }
}

// Precondition for lane s: nf->init[s] >= 0 (start GPU available); callers test it once per node.
__device__ __forceinline__ void greedy_chain_fast(const NodeFast* nf, int s, int k, int32_t& key, uint32_t& pay) {
    // a compare tree (two or three tests per size) rather than a switch: the switch became an indexed constant load
    // plus an indirect branch, with the load's latency in front of every short chain
    if (k < 5) {
        if (k < 3) { if (k == 2) greedy_chain_fast_k<2>(nf, s, key, pay); else { key = 0; pay = 0; } }
        else if (k == 3) greedy_chain_fast_k<3>(nf, s, key, pay);
        else greedy_chain_fast_k<4>(nf, s, key, pay);
    } else if (k < 7) {
        if (k == 5) greedy_chain_fast_k<5>(nf, s, key, pay); else greedy_chain_fast_k<6>(nf, s, key, pay);
    } else if (k == 7) greedy_chain_fast_k<7>(nf, s, key, pay);
    else if (k == 8) greedy_chain_fast_k<8>(nf, s, key, pay);
    else { key = 0; pay = 0; }
}

// estimateBandwidth, scheduler.go:656-692: mean over pairs i<j (SelectedGPUs order) of GPU
// i's link entry to GPU j when one exists (IsActive is NOT checked there), else GPU i's
// PCIe bandwidth.  Ordered FP64 sum.
__device__ inline double est_bandwidth(const DevNode* rec, uint32_t order, int n) {
    if (n <= 1) return 0.0;
    double total = 0.0; int pairs = 0;
    for (int i = 0; i < n; ++i) {
        int a = (order >> (3 * i)) & 7;
        for (int j = i + 1; j < n; ++j) {
            int b = (order >> (3 * j)) & 7;
            double x = (rec->present[a] >> b & 1) ? rec->bw_in[b][bw_slot(b, a)] : rec->pcie_bw[a];
            total = __dadd_rn(total, x); ++pairs;
        }
    }
    return pairs ? __ddiv_rn(total, (double)pairs) : 0.0;
}

// One greedy chain of findBestNVLinkGroup (scheduler.go:389-426) from start GPU `s`.
// agg[c] is carried across steps: the reference recomputes the candidate's sum over the
// members in join order from 0.0 at every step, which is exactly this running prefix sum
// (same operands, same order), so the results are bit-identical.
__device__ __forceinline__ void greedy_chain(const DevNode* rec, const NodeCtx& nc, int s, int k,
                                             bool& ok, double& total, uint32_t& order, uint32_t& members) {
    ok = (nc.avail >> s) & 1;
    total = 0.0; order = (uint32_t)s; members = 1u << s;
    if (!ok) return;
    double agg[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) agg[c] = 0.0;
    uint32_t conn = nc.avail & ~members;
    int m = s;
    for (int step = 1; step < k; ++step) {
        conn &= (uint32_t)(nc.colact >> (8 * m)) & 0xffu;
        const double2* row = reinterpret_cast<const double2*>(rec->bw_in[m]);
        int f = m >> 1;
        double best = 0.0; int bc = -1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            double2 v = row[q ^ f];
            agg[2 * q] = __dadd_rn(agg[2 * q], v.x);
            agg[2 * q + 1] = __dadd_rn(agg[2 * q + 1], v.y);
            if ((conn >> (2 * q) & 1) && agg[2 * q] > best) { best = agg[2 * q]; bc = 2 * q; }
            if ((conn >> (2 * q + 1) & 1) && agg[2 * q + 1] > best) { best = agg[2 * q + 1]; bc = 2 * q + 1; }
        }
        if (bc < 0) { ok = false; return; }
        members |= 1u << bc; conn &= ~(1u << bc);
        order |= (uint32_t)bc << (3 * step);
        total = __dadd_rn(total, best);
        m = bc;
    }
}

// EXTENSION E1: lane j scores the subsets whose mask value is congruent to j mod 8.
__device__ inline void exhaustive_lane(const DevNode* rec, const NodeCtx& nc, int j, int k, double& best, uint32_t& bestmask) {
    best = 0.0; bestmask = 0;
    for (uint32_t sub = (uint32_t)j; sub < 256u; sub += 8u) {
        if ((sub & ~nc.avail) || __popc(sub) != k) continue;
        bool ok = true;
        for (int m = 0; m < 8; ++m)
            if ((sub >> m & 1) && (sub & ~(1u << m) & ~(uint32_t)((nc.colact >> (8 * m)) & 0xff))) ok = false;
        if (!ok) continue;
        double total = 0.0; bool first = true;
        for (int jj = 0; jj < 8; ++jj) {
            if (!(sub >> jj & 1)) continue;
            double agg = 0.0;
            for (int i = 0; i < jj; ++i)
                if (sub >> i & 1) agg = __dadd_rn(agg, rec->bw_in[i][bw_slot(i, jj)]);
            if (!first) total = __dadd_rn(total, agg);
            first = false;
        }
        if (total > best) { best = total; bestmask = sub; }
    }
}

__device__ __forceinline__ int nvlink_score(double total, int k, double norm) {   // scheduler.go:368-370
    double maxbw = __dmul_rn(norm, (double)(k * (k - 1) / 2));
    double ratio = __ddiv_rn(total, maxbw);
    return __double2int_rz(__dadd_rn(50.0, __dmul_rn(50.0, ratio)));
}

// ---- one evaluation, in three parts ----------------------------------------------------
// eval_filter : isNodeEligible (Filter)                                   scheduler.go:206-241
// NVLink group search (greedy chains / exhaustive), only for NVLink preferences with k >= 2
// eval_finish : the rest of calculateTopologyScore + resource/balance/total scheduler.go:244-333, 516-578
// `evaluate` strings them together for one pod per 8-lane group; the scoring kernel's octet
// path calls the parts itself so that each lane finishes a different pod.

__device__ __forceinline__ int clamp_k(int gpu_count) {
    // k > 8 can never be eligible (A <= 8); clamp so no loop depends on a huge count
    return gpu_count > 9 ? 9 : gpu_count;
}
__device__ __forceinline__ bool is_nvlink_pref(const kgwe_pod_req& pod) {
    return pod.topology_pref == KGWE_TOPO_NVLINK_OPTIMAL || pod.topology_pref == KGWE_TOPO_NVLINK_REQUIRED;
}

__device__ __forceinline__ bool eval_filter(const DevNode* rec, const NodeCtx& nc, const kgwe_pod_req& pod, int k) {
    bool eligible = nc.valid && k >= 0 && nc.A >= k;
    if (pod.arch != KGWE_ARCH_ANY) {
        uint32_t match = 0;
        for (int g = 0; g < 8; ++g) if ((uint8_t)(nc.arch8 >> (8 * g)) == pod.arch) match |= 1u << g;
        eligible = eligible && (match & nc.avail);
    }
    if (pod.mig_profile != KGWE_MIG_NONE) {
        int have = (pod.mig_profile >= 1 && pod.mig_profile <= KGWE_MIG_PROFILES && nc.valid) ? rec->mig_free[pod.mig_profile - 1] : 0;
        eligible = eligible && have >= pod.mig_count;
    }
    return eligible;
}

// For NVLink preferences with k >= 2 the caller passes the group search's outcome:
// found / win_total / sel / order (SelectedGPUs) and the candidate count.
__device__ __forceinline__ void eval_finish(const DevNode* rec, const NodeCtx& nc, const kgwe_pod_req& pod, const Params& P,
                                            bool hinted, bool eligible, int k,
                                            bool found, double win_total, uint32_t sel, uint32_t order, uint32_t candidates,
                                            EvalOut& o, int topo_known = -1) {
    o.total = o.topo = o.res = o.bal = 0; o.mask = o.order = o.flags = 0; o.candidates = 1;
    if (!eligible) return;
    int topo = 0; int nsel = 0; bool fallback = false;
    if (k <= 1) {
        sel = 0; order = 0;
        if (nc.A > 0) { topo = 100; first_k_avail(nc, 1, sel, order, nsel); }
    } else if (is_nvlink_pref(pod)) {
        o.candidates = candidates;
        if (!found) {                    // no fully connected group: scheduler.go:358-365
            topo = 30; fallback = true; first_k_avail(nc, k, sel, order, nsel);
        } else {
            topo = topo_known >= 0 ? topo_known : nvlink_score(win_total, k, P.norm);
            nsel = k;
        }
    } else if (pod.topology_pref == KGWE_TOPO_SAME_NUMA || pod.topology_pref == KGWE_TOPO_SAME_PCIE_SWITCH) {
        const bool numa = pod.topology_pref == KGWE_TOPO_SAME_NUMA;
        const uint64_t pick = numa ? nc.numa_pick : nc.pcie_pick;
        const uint32_t gorder = numa ? nc.numa_order : nc.pcie_order;
        const uint32_t grp = (uint32_t)((pick >> (8 * (k - 1))) & 0xff);      // 2 <= k <= A <= 8 here
        sel = 0; order = 0;
        if (grp) { topo = numa ? 90 : 80; take_first_k(grp, k, sel, order); nsel = k; }
        else {
            topo = numa ? 50 : 40; fallback = true;
            order = gorder & ((k >= 8) ? 0xffffffu : ((1u << (3 * k)) - 1u));
            sel = mask_of_order(order, k); nsel = k;
        }
    } else {
        topo = 50;
        first_k_avail(nc, k, sel, order, nsel);
    }
    // --- calculateResourceScore, scheduler.go:516-553 ---
    int res = 0;
    if (nc.A > 0) {
        const uint64_t required = (uint64_t)(uint32_t)pod.min_memory_gb * 1073741824ull * (uint64_t)(uint32_t)k;
        res = 50 + (nc.sum_free >= required * 2 ? 25 : (nc.sum_free >= required ? 10 : 0)) + nc.util_bonus;
        res = res > 100 ? 100 : res;
    }
    // --- weighted total, scheduler.go:282-286 ---
    int total = (topo * P.wt + res * P.wr + nc.balance * P.wb) / 100;
    uint32_t flags = KGWE_F_ELIGIBLE;
    if (hinted) { total += P.bonus; flags |= KGWE_F_HINT; }
    bool kept = total > 0;                                   // scheduler.go:196
    if (fallback) flags |= KGWE_F_FALLBACK;
    // EXTENSION E3: minimum inter-GPU bandwidth (field declared at types.go:44, unused by the reference)
    if (kept && pod.min_bandwidth_gbps > 0.0 && k >= 2 && est_bandwidth(rec, order, nsel) < pod.min_bandwidth_gbps) {
        kept = false; flags |= KGWE_F_MINBW;
    }
    if (kept) flags |= KGWE_F_KEPT;
    o.total = total; o.topo = topo; o.res = res; o.bal = nc.balance;
    o.mask = sel; o.order = order | ((uint32_t)nsel << 24); o.flags = flags;
}

#ifdef __CUDACC__      // warp-collective code (8 lanes per evaluation): K1b, K3
// Winner of the 8 integer chains of a group: key = (total << 3) | (7 - start).  `gmask` names the
// 8 lanes of the group (or the whole warp when all four groups are known to be here together).
__device__ __forceinline__ void fast_group_winner(int32_t& key, uint32_t& pay, uint32_t gmask) {
    // (REDUX.MAX was measured slower than three shuffle+max steps here: 5.32 vs 4.80 ms per 1e8 evaluations)
#pragma unroll
    for (int off = 4; off >= 1; off >>= 1) key = max(key, __shfl_xor_sync(gmask, key, off, 8));
    pay = __shfl_sync(gmask, pay, ~key & 7, 8);       // 7 - (key & 7) = the winning start lane
}

// isNodeEligible + scoreNode for one (pod,node), executed by the 8 lanes of a group
// (lane j = `lane & 7`); the result is identical on all 8 lanes.  The 8 lanes of a group must call
// it together; its shuffles name only the group's own lanes, so the four groups of a warp may be
// on different pods (and so on different branches).
template <bool FAST>
__device__ __forceinline__ void evaluate(const DevNode* rec, const NodeCtx& nc, const kgwe_pod_req& pod, const Params& P,
                                         bool hinted, int lane, EvalOut& o, const NodeFast* nf = nullptr, bool warp_fast = false) {
    const int j = lane & 7;
    const uint32_t gmask = 0xffu << (lane & 24);
    const int k = clamp_k(pod.gpu_count);
    const bool eligible = eval_filter(rec, nc, pod, k);
    bool found = false; double win_total = 0.0; uint32_t sel = 0, order = 0, candidates = 1;
    if (k >= 2 && is_nvlink_pref(pod)) {
        // scoreNVLinkTopology, scheduler.go:336-373 (warp-uniform branch: one pod per warp pass)
        const bool exhaustive = pod.flags & KGWE_POD_EXHAUSTIVE;
        if (FAST && warp_fast && !exhaustive) {
            // all four nodes of this warp have whole-number bandwidths: integer chains, 32-bit winner
            int32_t key = 0; uint32_t pay = 0;
            if (eligible && nf->init[j] >= 0) greedy_chain_fast(nf, j, k, key, pay);
            fast_group_winner(key, pay, gmask);
            found = key > 0; win_total = (double)(key >> 3);
            order = pay; sel = mask_of_order(order, k);
            candidates = (uint32_t)nc.A;
        } else {
            uint64_t key; uint32_t pay;
            if (!exhaustive) {
                bool ok = false; double total = 0.0; uint32_t ord = 0, mem = 0;
                if (eligible) greedy_chain(rec, nc, j, k, ok, total, ord, mem);
                key = (ok && total > 0.0) ? (uint64_t)__double_as_longlong(total) : 0ull;
                pay = ord | (mem << 24);
                candidates = (uint32_t)nc.A;
            } else {
                double best = 0.0; uint32_t bm = 0;
                if (eligible) exhaustive_lane(rec, nc, j, k, best, bm);
                key = best > 0.0 ? (uint64_t)__double_as_longlong(best) : 0ull;
                pay = bm;
                uint32_t c = 1; for (int i = 1; i <= k; ++i) c = c * (uint32_t)(nc.A - k + i) / (uint32_t)i;
                candidates = nc.A >= k ? c : 0;
            }
            // winner over the 8 lanes: strictly largest total, then lowest start index / lowest mask
            // (greedy: first start in available order, scheduler.go:428-431)
            uint32_t tie = exhaustive ? pay : (uint32_t)j;
#pragma unroll
            for (int off = 4; off >= 1; off >>= 1) {
                uint64_t okey = __shfl_xor_sync(gmask, key, off, 8);
                uint32_t otie = __shfl_xor_sync(gmask, tie, off, 8);
                uint32_t opay = __shfl_xor_sync(gmask, pay, off, 8);
                if (okey > key || (okey == key && otie < tie)) { key = okey; tie = otie; pay = opay; }
            }
            found = key != 0ull; win_total = __longlong_as_double((long long)key);
            if (exhaustive) take_first_k(pay, k, sel, order);
            else { order = pay & 0xffffffu; sel = pay >> 24; }
        }
    }
    eval_finish(rec, nc, pod, P, hinted, eligible, k, found, win_total, sel, order, candidates, o);
}

#endif  // __CUDACC__

__device__ __forceinline__ kgwe_node_score to_record(const EvalOut& o) {
    kgwe_node_score r;
    r.total = o.total; r.topology = (int16_t)o.topo; r.resource = (int16_t)o.res; r.balance = (int16_t)o.bal;
    r.gpu_mask = (uint8_t)o.mask; r.flags = (uint8_t)o.flags; r.order = o.order;
    return r;
}

// Ranking key: TotalScore descending, then node index ascending (canonical tie-break).
// Only kept evaluations rank (total > 0), so 0 means "none".
__device__ __forceinline__ uint64_t rank_key(int32_t total, uint32_t global_node) {
    return ((uint64_t)(uint32_t)total << 32) | (uint64_t)(0xFFFFFFFFu - global_node);
}
__device__ __forceinline__ uint32_t key_node(uint64_t key) { return 0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFu); }

}  // namespace kgwe
