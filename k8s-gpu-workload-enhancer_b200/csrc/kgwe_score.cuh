// kgwe_score.cuh — K1, the stateless Filter + Prioritize kernel (scoreNodes for a batch of pods,
// reference src/scheduler/scheduler.go:182-203 and everything it calls).
//
// grid = node tiles (K1_TILE nodes) x pod chunks; ONE LANE = ONE NODE (kgwe_lane.cuh).
//   set-up   the CTA decodes its pod chunk into 16-byte PodLite records in shared memory; every lane reads
//            its node's 768-byte record with 128-bit loads, derives the pod-independent state into registers
//            (LaneNode) and its tables into [entry][lane] columns of shared memory (conflict-free for any
//            per-lane entry): NVLink score per k of a uniform node, NUMA / PCIe group tables, MIG counts,
//            and - kernel variants - the int32 / float64 link matrix, the E1 subset keys, memoised winners.
//   loop     each warp walks the pods of the chunk for its 32 nodes: lane_evaluate (no cross-lane work),
//            one 16-byte NodeScore store per lane (a warp writes 512 contiguous bytes of the pod's row),
//            one REDUX.MAX + one shared-memory atomicMax per warp for the per-pod arg-max
//            (TotalScore << 7 | 127 - node-in-tile), flushed with one global 64-bit atomicMax per pod per CTA.
// Variants: GENERAL adds the matrix columns (images that hold a node whose active links are not all equal);
// TABLES adds the E1 subset-key columns (implies GENERAL).  The search mode (all starts / first completed walk /
// memoised winners) is a launch argument.
#pragma once
#include "kgwe_lane.cuh"

namespace kgwe {

constexpr int K1_THREADS = 128;                // 4 warps
constexpr int K1_TILE = K1_THREADS;            // nodes per CTA: one per lane
constexpr int MAX_PODS_PER_CHUNK = 1024;       // size of the per-CTA PodLite / arg-max tables

// Optional fused finalisation for the latency path: out != nullptr enables it.
// Up to 4 pod descriptors can travel inside the kernel parameters (n_inline > 0): the kernel then never
// reads the descriptors over PCIe from pinned host memory, which removes dependent host-memory round trips
// from a single-pod call.
struct FusedTail { kgwe_placement* out; unsigned int* done; unsigned int* flag; unsigned int seq; unsigned int n_inline; kgwe_pod_req inline_pods[4]; };

struct K1Layout { uint32_t pods, keys, topo, grp, mig, memo, mat, subset, slot, total; };
// mat_units: 16-byte units per lane of the matrix columns: 0 (lean variant), 16 (int32 rows), 32 (int32 or float64 rows)
// groups: warp groups of K1_TILE lanes per CTA that share the node tables and split the pod chunk (1 or 2)
__host__ __device__ inline K1Layout k1_layout(uint32_t ppc, uint32_t mat_units, bool tables, bool memo, uint32_t groups = 1) {
    K1Layout L; uint32_t o = 0;
    L.pods = o; o += (ppc + 1u) * 16u;                              // + one slot: the loop prefetches pod i + 1
    L.keys = o; o += ((((ppc + groups - 1u) / groups) * 4u * (K1_THREADS / 32) * groups) + 15u) & ~15u;    // one arg-max row per warp over its group's share of the chunk (plain stores, no atomics)
    L.mat = o; o += mat_units * K1_TILE * 16u;
    L.memo = o; o += memo ? 7u * K1_TILE * 8u : 0u;
    L.topo = o; o += 7u * K1_TILE * 4u;
    L.grp = o; o += 6u * K1_TILE * 4u;
    L.mig = o; o += 3u * K1_TILE * 4u;
    L.subset = o; o += tables ? (uint32_t)SUBSET_UNITS * K1_TILE * 16u : 0u;
    L.slot = o; o += tables ? 512u : 0u;
    L.total = o;
    return L;
}

// The variants with big per-node tables (float64 matrix columns, subset keys) fit few CTAs per SM, i.e. few warps to hide the
// latency of their dependent chains.  They are launched with TWO warp groups of K1_TILE lanes per CTA: both groups own the same
// 128 nodes and share their tables (built once, by group 0), and each takes half of the CTA's pod chunk.
constexpr int K1_MAX_GROUPS = 2;
template <bool GENERAL, bool TABLES>
__global__ void __launch_bounds__(TABLES ? K1_THREADS * 4 : GENERAL ? K1_THREADS * K1_MAX_GROUPS : K1_THREADS, TABLES ? 1 : (GENERAL ? 2 : 5))
score_kernel(const DevNode* __restrict__ nodes, uint32_t n_nodes,
             const kgwe_pod_req* __restrict__ pods, uint32_t n_pods, uint32_t pods_per_chunk,
             const uint8_t* __restrict__ hint, uint32_t hint_row_bytes,
             kgwe_node_score* __restrict__ out, unsigned long long* __restrict__ best,
             Params P, unsigned long long* __restrict__ counters, uint32_t mode, uint32_t mat_units, FusedTail tail) {
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t n_groups = GENERAL ? (blockDim.x / K1_TILE) : 1u;
    const K1Layout L = k1_layout(pods_per_chunk, GENERAL ? mat_units : 0u, TABLES, mode == K1_MODE_MEMO, n_groups);
    PodLite* spod = reinterpret_cast<PodLite*>(smem + L.pods);
    uint32_t* keys = reinterpret_cast<uint32_t*>(smem + L.keys);
    uint16_t* slot_of = reinterpret_cast<uint16_t*>(smem + L.slot);             // TABLES only: entry of each mask

    const int tid = threadIdx.x, lane = tid & 31;
    const int ltid = GENERAL ? (tid & (K1_TILE - 1)) : tid;            // the lane's node within the tile
    const uint32_t grp = GENERAL ? (uint32_t)tid / K1_TILE : 0u;      // its warp group
    const uint32_t n_threads = GENERAL ? blockDim.x : (uint32_t)K1_THREADS;
    const uint32_t tile_first = blockIdx.x * K1_TILE;
    const uint32_t p_begin = blockIdx.y * pods_per_chunk;
    const uint32_t p_end = min(n_pods, p_begin + pods_per_chunk);
    const uint32_t np = p_end - p_begin;

    __shared__ kgwe_pod_req inline_copy[4];
    if (tid < (int)tail.n_inline) inline_copy[tid] = tail.inline_pods[tid];
    const kgwe_pod_req* spods = tail.n_inline ? inline_copy : nullptr;

    // ---- set-up 1: pod chunk -> PodLite, and what the chunk needs from the per-node tables ----
    __shared__ uint32_t cta_needs;
    if (tid == 0) cta_needs = 0u;
    __syncthreads();
    uint32_t needs = 0;
    for (uint32_t i = tid; i < np; i += n_threads) {
        const kgwe_pod_req pod = load_pod2(pods, spods, p_begin + i);
        const PodLite pl = pod_lite(pod);
        spod[i] = pl;
        if (i == np - 1) spod[np] = pl;
        needs |= pod_needs(pl);
    }
    if (TABLES)      // slot of mask m in (popcount, value) order
        for (int m = tid; m < 256; m += (int)n_threads) {
            int rank = 0;
            for (int q = 0; q < m; ++q) rank += __popc(q) == __popc(m);
            slot_of[m] = (uint16_t)(subset_off(__popc(m)) + rank);
        }
    needs = __reduce_or_sync(0xffffffffu, needs);
    if (lane == 0 && needs) atomicOr(&cta_needs, needs);
    __syncthreads();
    needs = cta_needs;
    const bool want_groups = (needs & NEED_GROUPS) != 0u;

    // ---- set-up 2: this lane's node ----
    const uint32_t node = tile_first + (uint32_t)ltid;
    const bool in_range = node < n_nodes;
    const DevNode* rec = nodes + (in_range ? node : 0);
    LaneNode ln;
    LaneTabs tb;
    tb.stride = K1_TILE;
    tb.topo = reinterpret_cast<int32_t*>(smem + L.topo) + ltid;
    tb.grp = reinterpret_cast<uint32_t*>(smem + L.grp) + ltid;
    tb.mig = reinterpret_cast<uint32_t*>(smem + L.mig) + ltid;
    tb.memo = reinterpret_cast<int2*>(smem + L.memo) + ltid;
    tb.mat = reinterpret_cast<int4*>(smem + L.mat) + ltid;
    tb.subset = reinterpret_cast<uint32_t*>(smem + L.subset) + 4 * ltid;
    {
        NodeCtx c;
        node_ctx_core(c, rec, in_range);
        if (want_groups) node_ctx_group_tables(c, rec);
        lane_node_init(ln, c, rec, GENERAL ? mat_units : 0u);
        if (grp == 0) lane_tables_build<GENERAL, TABLES>(ln, c, tb, rec, P.norm, mode, slot_of, needs);      // (the groups share the tables)
    }
    __syncthreads();

    // ---- the pod loop: every warp scores all pods of the chunk against its 32 nodes ----
    uint32_t cnt_scored = 0, cnt_resolved = 0;
    const bool has_hint = hint != nullptr, has_best = best != nullptr;
    const uint32_t do_store = (out != nullptr && in_range) ? 1u : 0u;
    // this group's part of the chunk
    const uint32_t sub = (np + n_groups - 1u) / n_groups;
    const uint32_t g_begin = min(np, grp * sub), g_end = min(np, g_begin + sub);
    int4* outp = reinterpret_cast<int4*>(out) + ((size_t)(p_begin + g_begin) * n_nodes + node);              // only dereferenced when do_store
    const uint8_t* hintp = hint + ((size_t)(p_begin + g_begin) * hint_row_bytes + (node >> 3));               // only dereferenced when has_hint && in_range
    const uint32_t key_low = (uint32_t)(K1_TILE - 1 - ltid);
    uint32_t pod_addr = (uint32_t)__cvta_generic_to_shared(spod) + g_begin * 16u;
    const uint32_t row_len = (pods_per_chunk + n_groups - 1u) / n_groups;                                      // a warp's row covers its group's share
    uint32_t key_addr = (uint32_t)__cvta_generic_to_shared(keys) + (uint32_t)(tid >> 5) * row_len * 4u;      // this warp's row, entry 0 = pod g_begin
    PodLite pd = spod[g_begin];
#pragma unroll 1
    for (uint32_t i = g_begin; i < g_end; ++i) {
        PodLite next;      // pod i + 1 (slot np repeats the last pod), straight from the shared window
        {
            uint32_t a0, a1, a2, a3;
            pod_addr += 16u;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3) : "r"(pod_addr));
            next.required = ((uint64_t)a1 << 32) | a0; next.mig_count = (int32_t)a2; next.pk = a3;
        }
        double min_bw = 0.0;
        if (pd.pk & PL_MINBW) min_bw = spods ? spods[p_begin + i].min_bandwidth_gbps : __ldg(&pods[p_begin + i].min_bandwidth_gbps);
        bool hinted = false;
        if (has_hint) { hinted = in_range && ((__ldg(hintp) >> (node & 7)) & 1); hintp += hint_row_bytes; }
        LaneResult r;
        lane_evaluate<GENERAL, TABLES>(ln, tb, rec, pd, min_bw, hinted, P, mode, r);
        cnt_scored += r.scored; cnt_resolved += r.candidates;
        if (do_store) *outp = make_int4((int)r.w0, (int)r.w1, (int)r.w2, (int)r.w3);
        outp += n_nodes;
        if (has_best) {
            // kept evaluations have TotalScore > 0, so key > 0 exactly when one of the 32 nodes is kept
            const uint32_t key = __reduce_max_sync(0xffffffffu, r.kept_total ? ((r.kept_total << 7) | key_low) : 0u);
            if (lane == 0) asm volatile("st.shared.u32 [%0], %1;" ::"r"(key_addr), "r"(key) : "memory");
            key_addr += 4u;
        }
        pd = next;
    }
    // counters: [1] candidates whose score was computed, [2] candidates by the reference's counting rule
    if (!in_range) { cnt_scored = 0; cnt_resolved = 0; }      // (lanes past the image counted 1 + 1 per pod: nothing of the image)
    cnt_scored = __reduce_add_sync(0xffffffffu, cnt_scored);
    cnt_resolved = __reduce_add_sync(0xffffffffu, cnt_resolved);
    if (lane == 0 && counters) {
        if (cnt_scored) atomicAdd(&counters[1], (unsigned long long)cnt_scored);
        if (cnt_resolved) atomicAdd(&counters[2], (unsigned long long)cnt_resolved);
    }
    if (best) {
        __syncthreads();
        for (uint32_t i = tid; i < np; i += n_threads) {
            const uint32_t g = i / sub, w0 = g * (K1_THREADS / 32), j = i - g * sub;      // the four warps of the group that scored pod i
            uint32_t kk = keys[w0 * row_len + j];
#pragma unroll
            for (int w = 1; w < K1_THREADS / 32; ++w) kk = max(kk, keys[(w0 + w) * row_len + j]);
            if (kk) atomicMax(&best[p_begin + i], rank_key((int32_t)(kk >> 7), P.node_base + tile_first + ((uint32_t)(K1_TILE - 1) - (kk & 127u))));
        }
    }
    if (tail.out) {
        // latency path (n_pods <= 32): the last CTA to finish turns the keys into placements in the
        // same launch and raises the host-visible flag (threadfence-reduction pattern)
        __shared__ unsigned int ticket;
        __threadfence();
        __syncthreads();
        if (tid == 0) ticket = atomicAdd(tail.done, 1u);
        __syncthreads();
        if (ticket == gridDim.x * gridDim.y - 1) {
            __threadfence();
            for (uint32_t p0 = 0; p0 < n_pods; p0 += K1_THREADS / 8) {
                const uint32_t p = p0 + (uint32_t)(tid >> 3);
                finalize_one(nodes, n_nodes, pods, p, p < n_pods, hint, hint_row_bytes, best, tail.out, P, lane, spods);
            }
            __threadfence_system();
            __syncthreads();
            if (tid == 0) { *tail.done = 0u; *reinterpret_cast<volatile unsigned int*>(tail.flag) = tail.seq; }
        }
    }
}

}  // namespace kgwe
