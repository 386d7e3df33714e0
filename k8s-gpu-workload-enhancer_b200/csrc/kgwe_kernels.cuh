// kgwe_kernels.cuh — the engine's kernels (sm_100a).
//
//  score_kernel      K1  stateless Filter+Prioritize, one lane per node: kgwe_score.cuh / kgwe_lane.cuh.
//  finalize_top1     K1b turns the arg-max keys into placements (+ EstimatedBandwidth).
//  topk_select       K2  ranked top-k per pod from the score matrix (canonical order).
//  topk_merge        K2b merges per-shard ranked lists (multi-GPU, after the all-gather).
//  schedule_kernel   K3  stateful queue: persistent cooperative kernel, one grid-wide arg-max
//                        and an owner-side commit per pod.
#pragma once
#include <cooperative_groups.h>

#include "kgwe_device.cuh"

namespace kgwe {
namespace cg = cooperative_groups;

constexpr int TILE_NODES = 64;                 // K3: nodes per CTA tile
constexpr int SCORE_THREADS = TILE_NODES * 8;  // K3: 8 lanes per node

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(phase) : "memory");
    }
}

// Arrival counter of the queue kernels' grid-wide step: release on the increment, acquire on the poll,
// instead of two full __threadfence()s around relaxed accesses.
__device__ __forceinline__ void arrive_release(unsigned int* ctr) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
}
__device__ __forceinline__ unsigned int load_acquire(const unsigned int* ctr) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
    return v;
}

__device__ __forceinline__ kgwe_pod_req load_pod(const kgwe_pod_req* pods, uint32_t p) {
    const uint4* src = reinterpret_cast<const uint4*>(pods + p);
    union { uint4 v[2]; kgwe_pod_req r; } u;
    u.v[0] = __ldg(src); u.v[1] = __ldg(src + 1);
    return u.r;
}

// Pod descriptor p from the shared-memory copy of the inline descriptors when there is one, else from global memory.
__device__ __forceinline__ kgwe_pod_req load_pod2(const kgwe_pod_req* pods, const kgwe_pod_req* spods, uint32_t p) {
    return spods ? spods[p] : load_pod(pods, p);
}

__device__ __forceinline__ bool hint_bit(const uint8_t* hint, uint32_t row_bytes, uint32_t p, uint32_t n) {
    return hint && ((__ldg(hint + (size_t)p * row_bytes + (n >> 3)) >> (n & 7)) & 1);
}

__device__ __forceinline__ void store_placement(kgwe_placement* dst, uint32_t node, const EvalOut& o, double est) {
    kgwe_placement pl;
    pl.node_idx = node; pl.total = o.total; pl.topology = (int16_t)o.topo; pl.resource = (int16_t)o.res;
    pl.balance = (int16_t)o.bal; pl.gpu_mask = (uint8_t)o.mask; pl.flags = (uint8_t)o.flags; pl.order = o.order;
    pl.reserved = 0; pl.est_bandwidth_gbps = est;
    *dst = pl;
}
__device__ __forceinline__ void store_no_placement(kgwe_placement* dst) {
    kgwe_placement pl;
    pl.node_idx = KGWE_NO_NODE; pl.total = 0; pl.topology = pl.resource = pl.balance = 0; pl.gpu_mask = 0; pl.flags = 0;
    pl.order = 0; pl.reserved = 0; pl.est_bandwidth_gbps = 0.0;
    *dst = pl;
}

// K1b: one 8-lane group per pod.  Turns the arg-max key into a placement: the winner's NodeScore comes from the
// score matrix when the call produced one (scores != nullptr), else the winning (pod,node) is evaluated again; adds
// EstimatedBandwidth, and clears the key slot for the next call.
__device__ __forceinline__ void finalize_one(const DevNode* __restrict__ nodes, uint32_t n_nodes, const kgwe_pod_req* __restrict__ pods,
                                             uint32_t p, bool live, const uint8_t* __restrict__ hint, uint32_t hint_row_bytes,
                                             unsigned long long* best, kgwe_placement* out, const Params& P, int lane,
                                             const kgwe_pod_req* spods = nullptr, const kgwe_node_score* scores = nullptr) {
    unsigned long long key = live ? *reinterpret_cast<volatile unsigned long long*>(&best[p]) : 0ull;
    const uint32_t node = key ? key_node(key) - P.node_base : 0;
    const DevNode* rec = nodes + node;
    EvalOut o;
    if (scores) {            // (uniform over the launch)
        o.total = o.topo = o.res = o.bal = 0; o.mask = o.order = o.flags = 0; o.candidates = 0;
        if (live && key != 0ull && node < n_nodes) {
            const int4 v = *reinterpret_cast<const int4*>(scores + (size_t)p * n_nodes + node);
            o.total = v.x; o.topo = (int)(int16_t)(v.y & 0xffff); o.res = (int)(int16_t)((uint32_t)v.y >> 16);
            o.bal = (int)(int16_t)(v.z & 0xffff); o.mask = ((uint32_t)v.z >> 16) & 0xffu; o.flags = (uint32_t)v.z >> 24; o.order = (uint32_t)v.w;
        }
    } else {
        kgwe_pod_req pod;
        if (live) pod = load_pod2(pods, spods, p); else { pod = kgwe_pod_req{}; }
        NodeCtx nc;
        node_ctx_core(nc, rec, live && key != 0ull && node < n_nodes);
        if (pod.topology_pref == KGWE_TOPO_SAME_NUMA || pod.topology_pref == KGWE_TOPO_SAME_PCIE_SWITCH) node_ctx_group_tables(nc, rec);
        evaluate<false>(rec, nc, pod, P, live && hint_bit(hint, hint_row_bytes, p, node), lane, o);
    }
    if (live && (lane & 7) == 0) {
        if (key == 0ull) store_no_placement(out + p);
        else store_placement(out + p, P.node_base + node, o, est_bandwidth(rec, o.order, (int)((o.order >> 24) & 15)));
        best[p] = 0ull;
    }
}

}  // namespace kgwe
#include "kgwe_score.cuh"      // K1 (needs load_pod2 / finalize_one above)
namespace kgwe {

// K1b with the exchange fused in: the placement of (this rank, pod p) is stored into slot `my_rank`
// of every rank's gather buffer over peer-mapped memory (NVLink P2P stores), replacing the all-gather.
// With flags (kgwe_score_merge_p2p): the last CTA to finish then tells every rank "my records of step `epoch` have
// landed" by writing the epoch into slot my_rank of that rank's flag array (system-scope fences on both sides of the
// CTA count order the record stores before the flag), so no separate cross-GPU barrier launch is needed.
struct PeerBufs { kgwe_placement* p[8]; };
struct PeerFlags { uint32_t* f[8]; };
__global__ void __launch_bounds__(256)
finalize_top1_p2p_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes, const kgwe_pod_req* __restrict__ pods, uint32_t n_pods,
                         const uint8_t* __restrict__ hint, uint32_t hint_row_bytes, const kgwe_node_score* __restrict__ scores,
                         unsigned long long* __restrict__ best, PeerBufs peers, uint32_t n_ranks, uint32_t my_rank, Params P,
                         PeerFlags flags, uint32_t epoch, unsigned int* __restrict__ ctas_done) {
    const uint32_t p = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const bool live = p < n_pods;
    // compute into this rank's own buffer first, then replicate the 32-byte record to the peers
    kgwe_placement* mine = peers.p[my_rank] + (size_t)my_rank * n_pods;
    finalize_one(nodes, n_nodes, pods, p, live, hint, hint_row_bytes, best, mine, P, threadIdx.x & 31, nullptr, scores);
    if (live && (threadIdx.x & 7) == 0) {
        const int4* src = reinterpret_cast<const int4*>(mine + p);
        const int4 a = src[0], b = src[1];
        for (uint32_t r = 0; r < n_ranks; ++r) {
            if (r == my_rank) continue;
            int4* dst = reinterpret_cast<int4*>(peers.p[r] + (size_t)my_rank * n_pods + p);
            dst[0] = a; dst[1] = b;
        }
    }
    if (ctas_done) {
        __shared__ unsigned int ticket;
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) ticket = atomicAdd(ctas_done, 1u);
        __syncthreads();
        if (ticket == gridDim.x - 1 && threadIdx.x < n_ranks) {
            __threadfence_system();
            if (threadIdx.x == 0) *ctas_done = 0u;
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flags.f[threadIdx.x] + my_rank), "r"(epoch) : "memory");
        }
    }
}

// The merge after that exchange: waits until every rank's flag in THIS rank's flag array has reached `epoch` (acquire
// loads; the flags only grow), then merges the R top-1 records of each pod (score desc, node index asc).  A rank that never
// shows up within ~4e9 cycles makes the kernel give up and raise my_flags[n_ranks] instead of spinning for ever.
__global__ void __launch_bounds__(128)
merge_top1_wait_kernel(const kgwe_placement* __restrict__ lists, uint32_t* my_flags, uint32_t n_ranks, uint32_t epoch, uint32_t n_pods,
                       kgwe_placement* __restrict__ out) {
    if (threadIdx.x < n_ranks) {
        const long long t0 = clock64();
        uint32_t v;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(my_flags + threadIdx.x) : "memory");
            if ((int32_t)(v - epoch) >= 0) break;
            if (clock64() - t0 > 4000000000ll) { my_flags[n_ranks] = 1u; break; }
        }
    }
    __syncthreads();
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    int br = -1; int32_t bt = 0; uint32_t bn = 0;
    for (uint32_t r = 0; r < n_ranks; ++r) {
        const kgwe_placement* c = lists + (size_t)r * n_pods + p;
        const uint32_t cn = c->node_idx;
        if (cn == KGWE_NO_NODE) continue;
        const int32_t ct = c->total;
        if (br < 0 || ct > bt || (ct == bt && cn < bn)) { br = (int)r; bt = ct; bn = cn; }
    }
    if (br < 0) store_no_placement(out + p);
    else {
        const int4* src = reinterpret_cast<const int4*>(lists + (size_t)br * n_pods + p);
        int4* dst = reinterpret_cast<int4*>(out + p);
        dst[0] = src[0]; dst[1] = src[1];
    }
}

__global__ void __launch_bounds__(256)
finalize_top1_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes, const kgwe_pod_req* __restrict__ pods, uint32_t n_pods,
                     const uint8_t* __restrict__ hint, uint32_t hint_row_bytes, const kgwe_node_score* __restrict__ scores,
                     unsigned long long* __restrict__ best, kgwe_placement* __restrict__ out, uint32_t out_stride, Params P) {
    const uint32_t p = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    finalize_one(nodes, n_nodes, pods, p, p < n_pods, hint, hint_row_bytes, best, out, P, threadIdx.x & 31, nullptr, scores);
}

// Empty cluster image: every requested placement is "no node" (nothing to score; Schedule then ends in "no eligible nodes found", scheduler.go:138-141).
__global__ void __launch_bounds__(256)
no_node_fill_kernel(kgwe_placement* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    kgwe_placement z{};
    z.node_idx = KGWE_NO_NODE;
    out[i] = z;
}

// K2: one CTA per pod over that pod's row of the score matrix.
__global__ void __launch_bounds__(256)
topk_select_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes, const kgwe_node_score* __restrict__ scores,
                   uint32_t topk, kgwe_placement* __restrict__ out, Params P) {
    __shared__ unsigned long long wmax[8];
    __shared__ unsigned long long sel[KGWE_MAX_TOPK];
    const uint32_t p = blockIdx.x;
    const kgwe_node_score* row = scores + (size_t)p * n_nodes;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long bound = ~0ull;
    for (uint32_t r = 0; r < topk; ++r) {
        unsigned long long local = 0ull;
        for (uint32_t n = tid; n < n_nodes; n += 256) {
            const int4 v = __ldg(reinterpret_cast<const int4*>(row + n));
            const uint32_t flags = ((uint32_t)v.z >> 24) & 0xff;
            if (flags & KGWE_F_KEPT) {
                const unsigned long long key = rank_key(v.x, P.node_base + n);
                if (key < bound && key > local) local = key;
            }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) { unsigned long long o2 = __shfl_xor_sync(0xffffffffu, local, off); local = o2 > local ? o2 : local; }
        if (lane == 0) wmax[warp] = local;
        __syncthreads();
        unsigned long long m = wmax[0];
#pragma unroll
        for (int w = 1; w < 8; ++w) m = wmax[w] > m ? wmax[w] : m;
        if (tid == 0) sel[r] = m;
        bound = m;
        __syncthreads();
        if (m == 0ull) { for (uint32_t r2 = r + 1 + tid; r2 < topk; r2 += 256) sel[r2] = 0ull; break; }
    }
    __syncthreads();
    if (tid < (int)topk) {
        const unsigned long long key = sel[tid];
        kgwe_placement* dst = out + (size_t)p * topk + tid;
        if (key == 0ull) store_no_placement(dst);
        else {
            const uint32_t n = key_node(key) - P.node_base;
            const kgwe_node_score s = row[n];
            EvalOut o; o.total = s.total; o.topo = s.topology; o.res = s.resource; o.bal = s.balance;
            o.mask = s.gpu_mask; o.flags = s.flags; o.order = s.order; o.candidates = 0;
            store_placement(dst, P.node_base + n, o, est_bandwidth(nodes + n, s.order, (int)((s.order >> 24) & 15)));
        }
    }
}

// K2b: thread per pod; lists[r][p][topk] are each ranked; emit the ranked union's first topk.
__global__ void __launch_bounds__(128)
topk_merge_kernel(const kgwe_placement* __restrict__ lists, uint32_t n_lists, uint32_t n_pods, uint32_t topk,
                  kgwe_placement* __restrict__ out) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    uint8_t head[32];
    for (uint32_t r = 0; r < n_lists && r < 32; ++r) head[r] = 0;
    for (uint32_t j = 0; j < topk; ++j) {
        int br = -1; int32_t bt = 0; uint32_t bn = 0;
        for (uint32_t r = 0; r < n_lists && r < 32; ++r) {
            if (head[r] >= topk) continue;
            const kgwe_placement* c = lists + ((size_t)r * n_pods + p) * topk + head[r];
            const uint32_t cn = c->node_idx;
            if (cn == KGWE_NO_NODE) continue;
            const int32_t ct = c->total;
            if (br < 0 || ct > bt || (ct == bt && cn < bn)) { br = (int)r; bt = ct; bn = cn; }
        }
        kgwe_placement* dst = out + (size_t)p * topk + j;
        if (br < 0) store_no_placement(dst);
        else { *dst = lists[((size_t)br * n_pods + p) * topk + head[br]]; head[br]++; }
    }
}

// ---------------------------------------------------------------------------------------
// K3.  Cooperative launch; grid * 64 lane groups; group g owns nodes g, g + G, g + 2G, ...
// Ownership is static, so a commit only has to be seen by the group that made it: ONE grid
// barrier per workload (for the arg-max), none for the commit.
__global__ void __launch_bounds__(SCORE_THREADS, 2)
schedule_kernel(DevNode* nodes, uint32_t n_nodes, const kgwe_pod_req* __restrict__ pods, uint32_t n_pods,
                const uint8_t* __restrict__ hint, uint32_t hint_row_bytes,
                unsigned long long* __restrict__ best, unsigned int* __restrict__ arrive, kgwe_placement* __restrict__ out,
                Params P, unsigned long long* __restrict__ counters) {
    __shared__ unsigned long long wmax[SCORE_THREADS / 32];
    __shared__ unsigned long long winner;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_groups = gridDim.x * (SCORE_THREADS / 8);
    const uint32_t g = blockIdx.x * (SCORE_THREADS / 8) + (tid >> 3);
    const uint32_t iters = (n_nodes + n_groups - 1) / n_groups;
    const bool single = iters <= 1;
    NodeCtx nc;
    if (single) node_ctx_init(nc, nodes + g, g < n_nodes);
    unsigned long long cand = 0, evals = 0;

    for (uint32_t p = 0; p < n_pods; ++p) {
        const kgwe_pod_req pod = load_pod(pods, p);
        unsigned long long mykey = 0ull; EvalOut keep; uint32_t keepnode = 0;
        keep.total = keep.topo = keep.res = keep.bal = 0; keep.mask = keep.order = keep.flags = 0; keep.candidates = 0;
        for (uint32_t it = 0; it < iters; ++it) {
            const uint32_t node = g + it * n_groups;
            const DevNode* rec = nodes + node;
            if (!single) node_ctx_init(nc, rec, node < n_nodes);
            EvalOut o;
            evaluate<false>(rec, nc, pod, P, node < n_nodes && hint_bit(hint, hint_row_bytes, p, node), lane, o);
            if (node < n_nodes) { cand += o.candidates; evals += 1; }
            const unsigned long long key = (o.flags & KGWE_F_KEPT) ? rank_key(o.total, P.node_base + node) : 0ull;
            if (key > mykey) { mykey = key; keep = o; keepnode = node; }
        }
        unsigned long long m = mykey;
#pragma unroll
        for (int off = 16; off >= 8; off >>= 1) { unsigned long long o2 = __shfl_xor_sync(0xffffffffu, m, off); m = o2 > m ? o2 : m; }
        if (lane == 0) wmax[warp] = m;
        __syncthreads();
        // grid-wide arg-max + barrier in one: every CTA folds its maximum into best[p], then bumps the
        // pod's arrival counter; the pod is decided when all CTAs have arrived (all CTAs are co-resident:
        // cooperative launch).  One counter per pod, so nothing has to be reset inside the kernel.
        if (tid == 0) {
            unsigned long long b = wmax[0];
            for (int w = 1; w < SCORE_THREADS / 32; ++w) b = wmax[w] > b ? wmax[w] : b;
            if (b) atomicMax(&best[p], b);
            arrive_release(&arrive[p]);                       // orders the max before the arrival (release)
            while (load_acquire(&arrive[p]) < gridDim.x) { }   // ... and the winner read after it (acquire)
            winner = *reinterpret_cast<volatile unsigned long long*>(&best[p]);
        }
        __syncthreads();
        const unsigned long long win = winner;
        if (win == 0ull) {
            if (g == 0 && (lane & 7) == 0) store_no_placement(out + p);      // "no eligible nodes found", scheduler.go:138-141
        } else if (win == mykey) {
            // this group owns the winner: tryScheduleOnNode + recordAllocation (scheduler.go:625-653, 695-707)
            DevNode* rec = nodes + keepnode;
            if ((lane & 7) == 0) {
                store_placement(out + p, P.node_base + keepnode, keep, 0.0);     // EstimatedBandwidth: est_bandwidth_kernel, after the queue
                rec->allocated_mask = (uint8_t)(rec->allocated_mask | keep.mask);
                rec->n_allocs = rec->n_allocs + 1;
            }
            __syncwarp(0xffu << (lane & 24));
            if (single) node_ctx_init(nc, rec, true);
        }
    }
    if ((lane & 7) != 0) { cand = 0; evals = 0; }
#pragma unroll
    for (int off = 16; off >= 8; off >>= 1) { cand += __shfl_xor_sync(0xffffffffu, cand, off); evals += __shfl_xor_sync(0xffffffffu, evals, off); }
    if (lane == 0 && counters) { if (evals) atomicAdd(&counters[0], evals); if (cand) { atomicAdd(&counters[1], cand); atomicAdd(&counters[2], cand); } }
}

// ---------------------------------------------------------------------------------------
// K3 tile form (used when every node has its own lane group: n_nodes <= grid * 64).  CTA b owns
// the node tile [64b, 64b+64) for the whole queue: the tile, its NodeCtx / NodeFast and a table of
// greedy winners per group size live in shared memory and are rebuilt by the owning group only
// when one of its commits changes the node.  Per workload: a filter + table look-up + finish per
// node, block max, then the same arg-max + arrival-counter barrier as above.  The next workload's
// descriptor is prefetched while the current one is decided.
__device__ __forceinline__ void group_winner_masked(int32_t& key, uint32_t& pay, uint32_t gmask) { fast_group_winner(key, pay, gmask); }

// (Re)build one node's tables by its 8 lanes.  `full` = first build (NVLink matrix rows too);
// after a commit only availability-dependent parts change.  NUMA / PCIe group tables are left
// stale (tables_valid = false) and filled on demand by ensure_group_tables.
__device__ __forceinline__ void node_tables_build(const DevNode* rec, bool in_range, bool full, NodeCtx* ctx, NodeFast* nf, int2* memo, int j, uint32_t gmask) {
    NodeCtx c;
    node_ctx_core(c, rec, in_range);
    if (c.integral) { if (full) node_fast_build(nf, rec, c, j); else nf->init[j] = (c.avail >> j & 1) ? (7 - j) : FAST_NEG; }
    else nf->init[j] = FAST_NEG;
    if (j == 0) *ctx = c;
    __syncwarp(gmask);
    int32_t sums[7]; int steps = -1; uint32_t order = 0;
    if (c.integral) greedy_chain_fast_all(nf, j, sums, steps, order);
#pragma unroll
    for (int kk = 2; kk <= 8; ++kk) {
        const int32_t total = steps >= kk - 1 ? (sums[kk - 2] >> FAST_SHIFT) : 0;
        int32_t key = total > 0 ? ((total << 3) | (7 - j)) : 0;
        uint32_t pay = order & ((kk >= 8) ? 0xffffffu : ((1u << (3 * kk)) - 1u));
        group_winner_masked(key, pay, gmask);
        if (j == 0) memo[kk - 1] = make_int2(key, (int)pay);
    }
    __syncwarp(gmask);
}

// NUMA / PCIe group tables of a node, computed the first time a workload asks for them after
// the node changed.  Group-uniform scalar code, no shuffles: safe under divergence.
__device__ __forceinline__ void ensure_group_tables(NodeCtx* ctx, const DevNode* rec, int j, uint32_t gmask) {
    if (ctx->tables_valid) return;
    NodeCtx c = *ctx;
    node_ctx_group_tables(c, rec);
    __syncwarp(gmask);
    if (j == 0) *ctx = c;
    __syncwarp(gmask);
}

__global__ void __launch_bounds__(SCORE_THREADS, 2)
schedule_tile_kernel(DevNode* nodes, uint32_t n_nodes, const kgwe_pod_req* __restrict__ pods, uint32_t n_pods,
                     const uint8_t* __restrict__ hint, uint32_t hint_row_bytes,
                     unsigned long long* __restrict__ best, unsigned int* __restrict__ arrive, kgwe_placement* __restrict__ out,
                     Params P, unsigned long long* __restrict__ counters, const uint32_t* __restrict__ gang_of) {
    extern __shared__ __align__(128) unsigned char smem[];
    DevNode* tile = reinterpret_cast<DevNode*>(smem);
    NodeFast* fast = reinterpret_cast<NodeFast*>(smem + TILE_NODES * sizeof(DevNode));
    NodeCtx* ctxs = reinterpret_cast<NodeCtx*>(smem + TILE_NODES * (sizeof(DevNode) + sizeof(NodeFast)));
    int2* memos = reinterpret_cast<int2*>(smem + TILE_NODES * (sizeof(DevNode) + sizeof(NodeFast) + sizeof(NodeCtx)));
    uint64_t* bar = reinterpret_cast<uint64_t*>(memos + TILE_NODES * 8);
    __shared__ unsigned long long wmax[SCORE_THREADS / 32];
    __shared__ unsigned long long winner;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, grp = tid >> 3, j = tid & 7;
    const uint32_t gmask = 0xffu << (lane & 24);
    const uint32_t tile_first = blockIdx.x * TILE_NODES;
    const uint32_t tile_count = tile_first < n_nodes ? min((uint32_t)TILE_NODES, n_nodes - tile_first) : 0;
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    if (tid == 0 && tile_count) {
        const uint32_t bytes = tile_count * (uint32_t)sizeof(DevNode);
        mbar_expect_tx(bar, bytes);
        tma_bulk_g2s(tile, nodes + tile_first, bytes, bar);
    }
    if (tile_count) mbar_wait(bar, 0);
    __syncthreads();

    const uint32_t node = tile_first + grp;
    const bool node_ok = grp < (int)tile_count;
    DevNode* rec = tile + grp;
    NodeFast* nf = fast + grp;
    int2* memo = memos + grp * 8;
    node_tables_build(rec, node_ok, true, ctxs + grp, nf, memo, j, gmask);
    const NodeCtx& nc = ctxs[grp];
    bool warp_fast = __all_sync(0xffffffffu, nc.integral || !nc.valid);
    unsigned long long cand = 0, evals = 0;
    kgwe_pod_req pod = load_pod(pods, 0);
    // E5 gangs (gang_of != nullptr): all-or-nothing per run of equal gang numbers.  Every thread sees every decision, so
    // "a member of the current gang found no node" is known grid-wide without any extra exchange; a group remembers what
    // its node committed for the current gang and takes it back when the gang's last member has been decided.
    bool gang_failed = false; uint32_t gang_mask = 0, gang_commits = 0;
    uint32_t gang_cur = gang_of ? __ldg(gang_of) : 0u;

    for (uint32_t p = 0; p < n_pods; ++p) {
        const kgwe_pod_req next = load_pod(pods, p + 1 < n_pods ? p + 1 : p);      // prefetch
        const uint32_t gang_next = (gang_of && p + 1 < n_pods) ? __ldg(gang_of + p + 1) : 0xFFFFFFFFu;
        const int k = clamp_k(pod.gpu_count);
        const bool hinted = node_ok && hint_bit(hint, hint_row_bytes, p, node);
        EvalOut o;
        if (pod.topology_pref == KGWE_TOPO_SAME_NUMA || pod.topology_pref == KGWE_TOPO_SAME_PCIE_SWITCH) ensure_group_tables(ctxs + grp, rec, j, gmask);
        if (warp_fast && !(pod.flags & KGWE_POD_EXHAUSTIVE)) {      // warp-uniform
            const bool eligible = eval_filter(rec, nc, pod, k);
            bool found = false; double win_total = 0.0; uint32_t sel = 0, ord = 0;
            if (eligible && k >= 2 && is_nvlink_pref(pod)) {
                const int2 w = memo[k - 1];
                found = w.x > 0; win_total = (double)(w.x >> 3); ord = (uint32_t)w.y; sel = found ? mask_of_order(ord, k) : 0u;
            }
            eval_finish(rec, nc, pod, P, hinted, eligible, k, found, win_total, sel, ord, (uint32_t)nc.A, o);
        } else {
            evaluate<false>(rec, nc, pod, P, hinted, lane, o);
        }
        if (node_ok && j == 0) { cand += o.candidates; evals += 1; }
        const unsigned long long mykey = (o.flags & KGWE_F_KEPT) ? rank_key(o.total, P.node_base + node) : 0ull;
        unsigned long long m = mykey;
#pragma unroll
        for (int off = 16; off >= 8; off >>= 1) { unsigned long long o2 = __shfl_xor_sync(0xffffffffu, m, off); m = o2 > m ? o2 : m; }
        if (lane == 0) wmax[warp] = m;
        __syncthreads();
        if (tid == 0) {
            unsigned long long b = wmax[0];
            for (int w = 1; w < SCORE_THREADS / 32; ++w) b = wmax[w] > b ? wmax[w] : b;
            if (b) atomicMax(&best[p], b);
            arrive_release(&arrive[p]);                       // orders the max before the arrival (release)
            while (load_acquire(&arrive[p]) < gridDim.x) { }   // ... and the winner read after it (acquire)
            winner = *reinterpret_cast<volatile unsigned long long*>(&best[p]);
        }
        __syncthreads();
        const unsigned long long win = winner;
        bool changed = false;
        if (win == 0ull) {
            if (blockIdx.x == 0 && tid == 0) store_no_placement(out + p);
            gang_failed = true;
        } else if (win == mykey) {
            // this group owns the winner: tryScheduleOnNode + recordAllocation (scheduler.go:625-653, 695-707)
            if (j == 0) {
                store_placement(out + p, P.node_base + node, o, 0.0);     // EstimatedBandwidth is filled in after the queue (est_bandwidth_kernel)
                rec->allocated_mask = (uint8_t)(rec->allocated_mask | o.mask);
                rec->n_allocs = rec->n_allocs + 1;
            }
            gang_mask |= o.mask; gang_commits += 1;
            changed = true;
        }
        if (gang_of && gang_next != gang_cur) {          // the gang's last member has been decided
            if (gang_failed && gang_commits) {            // take this node's commits for the gang back
                if (j == 0) {
                    rec->allocated_mask = (uint8_t)(rec->allocated_mask & ~gang_mask);
                    rec->n_allocs = rec->n_allocs - gang_commits;
                }
                changed = true;
            }
            gang_failed = false; gang_mask = 0; gang_commits = 0; gang_cur = gang_next;
        }
        if (changed) {
            if (j == 0) {
                nodes[node].allocated_mask = rec->allocated_mask;       // the image in HBM follows
                nodes[node].n_allocs = rec->n_allocs;
            }
            __syncwarp(gmask);
            node_tables_build(rec, true, false, ctxs + grp, nf, memo, j, gmask);
        }
        __syncwarp();
        pod = next;
    }
    if (j != 0) { cand = 0; evals = 0; }
#pragma unroll
    for (int off = 16; off >= 8; off >>= 1) { cand += __shfl_xor_sync(0xffffffffu, cand, off); evals += __shfl_xor_sync(0xffffffffu, evals, off); }
    // (winners come from the owner's per-size table: one candidate looked up per evaluation, A resolved)
    if (lane == 0 && counters) { if (evals) { atomicAdd(&counters[0], evals); atomicAdd(&counters[1], evals); } if (cand) atomicAdd(&counters[2], cand); }
}

// EstimatedBandwidth of the queue's decisions (scheduler.go:656-692), one thread per decision.  It
// depends only on link data that commits never change, so it is taken off the per-workload critical
// path of the queue kernels and computed for all decisions at once afterwards.
__global__ void __launch_bounds__(128)
est_bandwidth_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes, uint32_t node_base, kgwe_placement* __restrict__ pls, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t node = pls[i].node_idx;
    if (node == KGWE_NO_NODE || node - node_base >= n_nodes) return;
    const uint32_t order = pls[i].order;
    pls[i].est_bandwidth_gbps = est_bandwidth(nodes + (node - node_base), order, (int)((order >> 24) & 15));
}

// ---------------------------------------------------------------------------------------
// K5: DiscoveryService.GetTopologyHint (discovery.go:222-247, 378-488).  One thread per
// (request, node): a handful of mask operations; block max of score << 32 | ~node, one atomicMax
// per request per block; the winners are turned into hints by hint_finalize_kernel.
struct HintEval { int score; uint32_t mask, order; int n; };

__device__ __forceinline__ HintEval hint_eval(const DevNode* rec, bool in_range, const kgwe_hint_req& rq) {
    HintEval h{0, 0, 0, 0};
    if (!in_range || !(rec->flags & NODE_PRESENT)) return h;
    const uint32_t ng = rec->n_gpus > 8 ? 8 : rec->n_gpus;
    const uint32_t all = (1u << ng) - 1u;
    // getAvailableGPUs, discovery.go:436-460: healthy and (MIG enabled ? a free instance : GPUPercent < 90)
    uint32_t avail = 0;
    for (uint32_t g = 0; g < ng; ++g) {
        if (!(rec->healthy_mask >> g & 1)) continue;
        const bool ok = (rec->mig_enabled_mask >> g & 1) ? (rec->mig_has_free_mask >> g & 1) : (rec->util[g] < 90.0);
        if (ok) avail |= 1u << g;
    }
    const int count = rq.count;
    if ((int)__popc(avail) < count) return h;                                   // :380-382
    uint32_t filt = 0;                                                           // :385-391
    for (uint32_t g = 0; g < ng; ++g)
        if ((avail >> g & 1) && (long long)(rec->free_bytes[g] >> 30) >= (long long)rq.min_memory_gb) filt |= 1u << g;
    if ((int)__popc(filt) < count) return h;                                    // :393-395
    int score = 50; uint32_t mask = 0, order = 0; int n = 0;
    if (count > 1 && rq.prefer_nvlink) {                                         // :402-409 + findNVLinkGroups :462-488
        uint32_t m = filt;
        while (m) {
            const int g = __ffs(m) - 1; m &= m - 1;
            const uint32_t peers = rec->present[g] & filt & all & ~(1u << g);    // any link entry, active or not
            if ((int)__popc(peers) >= count - 1) {
                uint32_t s2, o2;
                take_first_k(peers, count - 1, s2, o2);
                mask = s2 | (1u << g); order = (uint32_t)g | (o2 << 3); n = count;
                score += 30;
                break;
            }
        }
    }
    if (n == 0) {                                                                // :411-415
        const int want = count < 0 ? 0 : count;
        take_first_k(filt, want > 8 ? 8 : want, mask, order); n = __popc(mask);
    }
    if (rq.prefer_same_numa) {                                                   // :418-420, areOnSameNUMA :490-504
        bool same = true; int first = -1000;
        for (uint32_t g = 0; g < ng; ++g) if (mask >> g & 1) { if (first == -1000) first = rec->numa[g]; else if (rec->numa[g] != first) same = false; }
        if (same) score += 10;
    }
    if (rq.arch != KGWE_ARCH_ANY)                                                // :423-431
        for (uint32_t g = 0; g < ng; ++g) if ((mask >> g & 1) && rec->arch[g] == rq.arch) score += 5;
    h.score = score; h.mask = mask; h.order = order | ((uint32_t)n << 24); h.n = n;
    return h;
}

__global__ void __launch_bounds__(256)
hint_score_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes, uint32_t node_base, const kgwe_hint_req* __restrict__ reqs,
                  unsigned long long* __restrict__ best) {
    __shared__ unsigned long long wmax[8];
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    const kgwe_hint_req rq = reqs[blockIdx.y];
    const HintEval h = hint_eval(nodes + node, node < n_nodes, rq);
    unsigned long long key = h.score > 0 ? rank_key(h.score, node_base + node) : 0ull;   // score > hint.Score(0), lowest index on ties
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) { const unsigned long long o = __shfl_xor_sync(0xffffffffu, key, off); key = o > key ? o : key; }
    if ((threadIdx.x & 31) == 0) wmax[threadIdx.x >> 5] = key;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long b = wmax[0];
        for (int w = 1; w < 8; ++w) b = wmax[w] > b ? wmax[w] : b;
        if (b) atomicMax(&best[blockIdx.y], b);
    }
}

__global__ void __launch_bounds__(128)
hint_finalize_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes, uint32_t node_base, const kgwe_hint_req* __restrict__ reqs, uint32_t n_reqs,
                     unsigned long long* __restrict__ best, kgwe_topology_hint* __restrict__ out) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reqs) return;
    const unsigned long long key = best[r];
    best[r] = 0ull;
    kgwe_topology_hint t;
    t.node_idx = KGWE_NO_NODE; t.score = 0; t.order = 0; t.gpu_mask = 0; t.pad[0] = t.pad[1] = t.pad[2] = 0; t.bandwidth_estimate = 0.0;
    if (key) {
        const uint32_t node = key_node(key) - node_base;
        const HintEval h = hint_eval(nodes + node, node < n_nodes, reqs[r]);
        t.node_idx = node_base + node; t.score = h.score; t.order = h.order; t.gpu_mask = (uint8_t)h.mask;
        t.bandwidth_estimate = est_bandwidth(nodes + node, h.order, h.n);          // discovery.go:506-541
    }
    out[r] = t;
}

// K6: the exporter's per-node topology gauge (prometheus_exporter.go:517-539).
__global__ void __launch_bounds__(256)
gauge_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes, int32_t* __restrict__ out, uint32_t n_out) {
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= n_out) return;
    int score = 0;
    if (node < n_nodes && (nodes[node].flags & NODE_PRESENT) && nodes[node].n_gpus > 0) {
        score = 50 + ((nodes[node].node_flags & KGWE_NODE_NVSWITCH) ? 30 : 0) + ((nodes[node].node_flags & KGWE_NODE_ANY_NVLINK) ? 20 : 0);
        score = score > 100 ? 100 : score;
    }
    out[node] = score;
}

// ---------------------------------------------------------------------------------------
// E4 side terms: one warp per placement.  Lanes split the 2^k subsets (bisection) and the
// (k-1)! cyclic orders (ring); warp-shuffle min / max at the end.  Cut sums are accumulated
// per lane in ascending (a, b) order, so the result does not depend on the lane split.
__device__ __forceinline__ double pair_w(const DevNode* rec, int a, int b) {
    const bool ab = rec->act_in[b] >> a & 1, ba = rec->act_in[a] >> b & 1;
    if (!ab || !ba) return 0.0;
    const double x = rec->bw_in[b][bw_slot(b, a)], y = rec->bw_in[a][bw_slot(a, b)];
    return x < y ? x : y;
}

__global__ void __launch_bounds__(128)
side_terms_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes, uint32_t node_base,
                  const kgwe_placement* __restrict__ pls, uint32_t n, kgwe_side_terms* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (i >= n) return;
    const kgwe_placement pl = pls[i];
    const uint32_t local = pl.node_idx - node_base;
    double bis = 0.0, ring = 0.0;
    int g[8]; int k = 0;
    if (pl.node_idx != KGWE_NO_NODE && local < n_nodes)
        for (int b = 0; b < 8; ++b) if (pl.gpu_mask >> b & 1) g[k++] = b;
    if (k >= 2) {
        const DevNode* rec = nodes + local;
        double w[8][8];
        for (int a = 0; a < k; ++a) for (int b = 0; b < k; ++b) w[a][b] = a == b ? 0.0 : pair_w(rec, g[a], g[b]);
        // bisection: subsets T of positions with |T| = k/2
        const int h = k / 2;
        double best = 1e300;
        for (uint32_t t = (uint32_t)lane; t < (1u << k); t += 32) {
            if (__popc(t) != h) continue;
            double cut = 0.0;
            for (int a = 0; a < k; ++a) if (t >> a & 1)
                for (int b = 0; b < k; ++b) if (!(t >> b & 1)) cut = __dadd_rn(cut, w[a][b]);
            best = cut < best ? cut : best;
        }
        // ring: position 0 fixed, the other k-1 permuted (factoradic decode of the order index)
        int nperm = 1; for (int f = 2; f < k; ++f) nperm *= f;
        double rbest = -1.0;
        for (int idx = lane; idx < nperm; idx += 32) {
            int rest[7]; for (int q = 0; q < k - 1; ++q) rest[q] = q + 1;
            int code = idx, prev = 0; double mn = 1e300;
            for (int q = k - 1; q >= 1; --q) {
                const int c = code % q; code /= q;
                const int cur = rest[c];
                for (int z = c; z < q - 1; ++z) rest[z] = rest[z + 1];
                const double e = w[prev][cur]; mn = e < mn ? e : mn;
                prev = cur;
            }
            const double e = w[prev][0]; mn = e < mn ? e : mn;
            rbest = mn > rbest ? mn : rbest;
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, off), orr = __shfl_xor_sync(0xffffffffu, rbest, off);
            best = ob < best ? ob : best; rbest = orr > rbest ? orr : rbest;
        }
        bis = best; ring = rbest;
    }
    if (lane == 0) { out[i].bisection_gbps = bis; out[i].ring_bottleneck_gbps = ring; }
}

}  // namespace kgwe
