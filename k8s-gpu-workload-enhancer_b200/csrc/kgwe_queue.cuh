// kgwe_queue.cuh — K3c, the stateful queue (Schedule for a queue of workloads with a commit after each one,
// reference src/scheduler/scheduler.go:114-179, 625-653, 695-707) inside ONE thread-block cluster.
//
// The queue is a dependent chain - workload p + 1 must see workload p's commit - so what bounds it is the latency
// of one grid-wide arg-max per workload, not throughput.  schedule_tile_kernel (kgwe_kernels.cuh) spreads the nodes
// over all SMs and pays two or three L2 round trips per workload for that arg-max.  Here the whole image sits in
// one cluster of up to 16 CTAs (ONE LANE = ONE NODE, state in registers and [entry][lane] shared-memory columns as
// in K1, kgwe_lane.cuh), and the arg-max goes through distributed shared memory:
//   every lane evaluates the workload on its node; REDUX.MAX per warp; one bar.sync; warp 0 folds the warp keys
//   and sends the CTA's key to EVERY CTA of the cluster as a remote atomic max (red.async on shared::cluster) that
//   also completes 4 bytes of the receiving CTA's mbarrier; one lane per warp waits for its own CTA's mbarrier
//   phase (all <= 16 keys have landed) and every thread reads the one word that now holds the winner's key.
//   No fence and no cluster-wide barrier inside the loop.
//   (the same wait CUTLASS uses for data that arrives with complete_tx: default acquire at CTA scope).
//   The lane that owns the winning node writes the decision and commits in the image; the state its node has after
//   the commit was worked out, while the keys were travelling, by the one lane of each CTA that could win.
// Keys are TotalScore << 14 | (16383 - node): largest score, lowest node index on ties (the canonical order).
// The greedy winners per group size are memoised per lane and filled lazily: a commit only invalidates the owner's
// memo, and the next evaluation that needs an entry runs the search of findBestNVLinkGroup (:376-435) for that size
// alone (on a uniform node the first completed walk, else all starts: every completed walk ties and the first wins).
// EstimatedBandwidth (:656-692) is filled in for all decisions by one CTA after the loop.
// Nodes of every link tier are served: the memo fill of a non-uniform node runs the float64 chains on the record
// itself (the lane variant without matrix columns), which a commit-bound queue can afford.
// Form limits (the engine falls back to schedule_tile_kernel otherwise): n_nodes <= cluster size x QC_MAX_THREADS
// (10,240 with 16 CTAs), TotalScore < 2^18.
#pragma once
#include "kgwe_kernels.cuh"
#include "kgwe_queue_state.cuh"

namespace kgwe {

constexpr int QC_MAX_THREADS = 640;        // 20 warps; 296 bytes of shared-memory columns per lane
constexpr int QC_SMALL_THREADS = 128;      // the form for images of up to 16 x 128 nodes: fewer warps per scheduler, shorter passes
constexpr int QC_MAX_CTAS = 16;            // non-portable cluster size, opt-in at launch
constexpr int QC_WINDOW = 256;             // pod descriptors staged per refill
constexpr uint32_t QC_NODE_BITS = 14;
constexpr uint32_t QC_MAX_NODES = 1u << QC_NODE_BITS;

struct QcLayout { uint32_t pods, minbw, gang, stat, topo, grp, mig, memo, wkeys, ckeys, bars, spec, total; };
__host__ __device__ inline QcLayout qc_layout(uint32_t threads) {
    QcLayout L; uint32_t o = 0;
    L.pods = o; o += QC_WINDOW * 16u;
    L.minbw = o; o += QC_WINDOW * 8u;
    L.gang = o; o += (QC_WINDOW + 1u) * 4u; o = (o + 15u) & ~15u;      // + the first gang number of the next window
    L.stat = o; o += (uint32_t)QS_ENTRIES * threads * 8u;
    L.memo = o; o += 7u * threads * 8u;
    L.topo = o; o += 7u * threads * 4u;
    L.grp = o; o += 6u * threads * 4u;
    L.mig = o; o += 3u * threads * 4u;
    L.wkeys = o; o += 2u * 32u * 4u;                                   // warp keys of even / odd workloads
    L.ckeys = o; o += 2u * QC_MAX_CTAS * 4u;
    L.bars = o; o += 4u * 8u;                                         // keys of even / odd workloads: inside the cluster, between the ranks
    L.spec = o; o += 16u * 4u;                                        // the CTA's best lane parks its would-be state here
    L.total = o;
    return L;
}

__device__ __forceinline__ uint32_t qc_cta_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t qc_map_to_cta(uint32_t smem_addr, uint32_t rank) {
    uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank)); return r;
}
// One 32-bit atomic max on a word of another CTA's shared memory that also signals that CTA's mbarrier (complete_tx of
// 4 bytes): the update is visible to whoever observes the phase of the mbarrier complete - no fence, no cluster-wide barrier.
__device__ __forceinline__ void qc_max_signal(uint32_t cluster_addr, uint32_t v, uint32_t cluster_mbar) {
    asm volatile("red.async.relaxed.cluster.mbarrier::complete_tx::bytes.shared::cluster.max.u32 [%0], %1, [%2];" ::"r"(cluster_addr), "r"(v), "r"(cluster_mbar) : "memory");
}
// One 32-bit store into another CTA's shared memory that completes 4 bytes of that CTA's mbarrier.
__device__ __forceinline__ void qc_store_signal(uint32_t cluster_addr, uint32_t v, uint32_t cluster_mbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.u32 [%0], %1, [%2];" ::"r"(cluster_addr), "r"(v), "r"(cluster_mbar) : "memory");
}

// Node-sharded queue (kgwe_schedule_queue_p2p): every rank runs this kernel on its shard of the cluster, one rank per
// GPU, and after the arg-max inside its cluster the ranks exchange ONE 64-bit word per workload through peer-mapped
// memory: tag << 40 | TotalScore << 22 | (2^22 - 1 - global node), written with a single 64-bit system-scope store into slot
// [sender][tag & 3] of every peer's receive array and polled there until the tag is the expected one (relaxed accesses: the
// word is the whole message, no other data rides on it).
// tag = the workload's sequence number (1 .. 2^24 - 1, running on across calls, the same on every rank): a slot only ever
// sees increasing tags, so a stale word cannot be taken for a fresh one, and a rank can be at most one workload ahead
// of a peer (it needs that peer's word to go on), so four slots per sender are more than enough.
struct QcPeers {
    unsigned long long* recv[8];   // recv[r]: the receive array of rank r ([n_ranks][4] 64-bit words, zero before the first call)
    uint32_t n_ranks, my_rank;
    uint32_t seq_base;             // sequence number of this call's first workload, 0 .. 2^24 - 2
    uint32_t pad;
};
constexpr uint32_t QC_PLACED_ELSEWHERE = 0x80000000u;   // what the cluster is told when another rank's node won
constexpr uint32_t QC_EXCHANGE_FAILED = 0xFFFFFFFFu;    // ... when a peer's word did not arrive in time
constexpr uint32_t QC_SEQ_MOD = 0xFFFFFFu;

__device__ __forceinline__ void qc_cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void qc_cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// Wait for `phase` of a local mbarrier whose transactions come from other CTAs.  Bounded: a protocol fault must not hang the
// GPU; the caller reports a trip through the watchdog counter.
__device__ __forceinline__ bool qc_wait_phase(uint32_t mbar, uint32_t phase) {
    for (uint32_t it = 0; it < (1u << 22); ++it) {
        uint32_t done;
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(mbar), "r"(phase) : "memory");
        if (done) return true;
    }
    return false;
}

// Lazily filled per-lane tables (by value: the lane state stays in registers across the call).
__device__ __noinline__ uint32_t qc_memo_fill(uint32_t avail, uint32_t rlo, uint32_t rhi, uint32_t tier, int2* memo, const int32_t* topo_tab, int S,
                                             const DevNode* rec, int k, double norm) {
    LaneNode ln; ln.avail = avail; ln.A = __popc(avail); ln.rlo = rlo; ln.rhi = rhi; ln.tier = tier; ln.valid = true;
    LaneTabs tb; tb.topo = const_cast<int32_t*>(topo_tab); tb.stride = S; tb.mat = nullptr;
    int topo = -1; uint32_t sel = 0, order = 0, sc = 0;
    lane_group_search<false>(ln, tb, rec, k, norm, K1_MODE_PRUNE, topo, sel, order, sc);      // (a uniform node's first completed walk is the winner)
    memo[(k - 2) * S] = make_int2(topo, (int)(order | (sel << 24)));
    return sc;
}
__device__ __noinline__ void qc_groups_fill(uint32_t avail, const uint64_t* st, uint32_t* grp, int S) {
    uint64_t np = 0, pp = 0; uint32_t no = 0, po = 0;
    if (((uint32_t)st[QS_MISC * S] >> 24) & 1u) {
        group_tables(st[QS_NUMA * S], true, avail, np, no);
        group_tables(st[QS_PCIE * S], false, avail, pp, po);
    }
    grp[0] = (uint32_t)np; grp[S] = (uint32_t)(np >> 32); grp[2 * S] = no;
    grp[3 * S] = (uint32_t)pp; grp[4 * S] = (uint32_t)(pp >> 32); grp[5 * S] = po;
}

template <int T, bool PEERS>
__global__ void __launch_bounds__(T, 1)
schedule_cluster_kernel(DevNode* nodes, uint32_t n_nodes, const kgwe_pod_req* __restrict__ pods, uint32_t n_pods,
                        const uint8_t* __restrict__ hint, uint32_t hint_row_bytes, kgwe_placement* __restrict__ out,
                        Params P, unsigned long long* __restrict__ counters, const uint32_t* __restrict__ gang_of, QcPeers PQ) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int n_warps = T >> 5;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t C = gridDim.x, rank = qc_cta_rank();
    const QcLayout L = qc_layout((uint32_t)T);
    PodLite* spod = reinterpret_cast<PodLite*>(smem + L.pods);
    double* sminbw = reinterpret_cast<double*>(smem + L.minbw);
    uint32_t* sgang = reinterpret_cast<uint32_t*>(smem + L.gang);
    uint32_t* wkeys = reinterpret_cast<uint32_t*>(smem + L.wkeys);
    uint32_t* ckeys = reinterpret_cast<uint32_t*>(smem + L.ckeys);
    volatile uint32_t* spec = reinterpret_cast<volatile uint32_t*>(smem + L.spec);
    uint64_t* st = reinterpret_cast<uint64_t*>(smem + L.stat) + tid;

    const uint32_t node = rank * (uint32_t)T + (uint32_t)tid;
    const bool in_range = node < n_nodes;
    DevNode* rec = nodes + (in_range ? node : 0);
    LaneNode ln;
    LaneTabs tb;
    tb.stride = T;
    tb.topo = reinterpret_cast<int32_t*>(smem + L.topo) + tid;
    tb.grp = reinterpret_cast<uint32_t*>(smem + L.grp) + tid;
    tb.mig = reinterpret_cast<uint32_t*>(smem + L.mig) + tid;
    tb.memo = reinterpret_cast<int2*>(smem + L.memo) + tid;
    tb.mat = nullptr; tb.subset = nullptr;
    if (tid < 2 * QC_MAX_CTAS) ckeys[tid] = 0u;
    const uint32_t bars = smem_u32(smem + L.bars);      // four mbarriers: the keys of even / odd workloads inside the cluster, the outcome of the exchange between the ranks (PEERS)
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8u));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 16u));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 24u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    uint32_t allocated = 0, n_allocs = 0;       // this node's ledger projection (scheduler.go:587-593, 562-566), mirrored in the image on every change
    {
        NodeCtx c;
        node_ctx_core(c, rec, in_range);
        lane_node_init(ln, c, rec, 0u);
        lane_tables_build<false, false>(ln, c, tb, rec, P.norm, K1_MODE_FULL, nullptr);      // NVLink score per k, MIG counts
        qc_static_build(st, T, rec, c, ln.tier);
        if (c.valid) { allocated = rec->allocated_mask; n_allocs = rec->n_allocs; }
    }
    qc_state(ln, st, T, allocated, n_allocs);
    uint32_t memo_ok = 0;        // bit k - 2: the memoised winner for group size k is current
    bool groups_ok = false;      // the NUMA / PCIe group tables are current
    __syncwarp();
    qc_cluster_arrive(); qc_cluster_wait();      // every CTA's key slots are initialised before anybody stores into them

    const uint32_t key_low = (QC_MAX_NODES - 1u) - node;
    const uint32_t ckeys_a = smem_u32(ckeys);
    uint32_t cnt_scored = 0, cnt_resolved = 0, cnt_evals = 0;
    // E5 gangs: see schedule_tile_kernel - every thread sees every decision; a lane remembers what its node committed
    // for the current gang and takes it back when the gang's last member has been decided and one of them found no node.
    bool gang_failed = false; uint32_t gang_mask = 0, gang_commits = 0;
    uint32_t hint_byte = (hint && in_range) ? __ldg(hint + (node >> 3)) : 0u;

    for (uint32_t p = 0; p < n_pods; ++p) {
        const uint32_t wi = p & (QC_WINDOW - 1);
        if (wi == 0) {      // stage the next window of pod descriptors
            __syncthreads();
            const uint32_t cnt = min((uint32_t)QC_WINDOW, n_pods - p);
            for (uint32_t i = tid; i <= cnt; i += T) {
                if (i < cnt) {
                    const kgwe_pod_req pod = load_pod(pods, p + i);
                    spod[i] = pod_lite(pod); sminbw[i] = pod.min_bandwidth_gbps;
                }
                sgang[i] = (gang_of && p + i < n_pods) ? __ldg(gang_of + p + i) : 0xFFFFFFFFu;
            }
            __syncthreads();
        }
        const PodLite pd = spod[wi];
        const bool hinted = (hint_byte >> (node & 7)) & 1u;
        if (hint && in_range && p + 1 < n_pods) hint_byte = __ldg(hint + (size_t)(p + 1) * hint_row_bytes + (node >> 3));      // next pod's bit, off the critical path
        double min_bw = 0.0;
        if (pd.pk & PL_MINBW) min_bw = sminbw[wi];
        const int k = (int)(pd.pk & 0xffu);
        if ((pd.pk & (PL_NVLINK | PL_EXH)) == PL_NVLINK) {      // warp-uniform
            if (ln.valid && ln.A >= k && !(memo_ok >> (k - 2) & 1u)) {
                cnt_scored += qc_memo_fill(ln.avail, ln.rlo, ln.rhi, ln.tier, tb.memo, tb.topo, T, rec, k, P.norm);
                memo_ok |= 1u << (k - 2);
            }
            __syncwarp();
        } else if (pd.pk & (PL_NUMA | PL_PCIE)) {
            if (ln.valid && ln.A >= k && !groups_ok) { qc_groups_fill(ln.avail, st, tb.grp, T); groups_ok = true; }
            __syncwarp();
        }
        LaneResult r;
        lane_evaluate<false, false>(ln, tb, rec, pd, min_bw, hinted, P, K1_MODE_MEMO, r);
        if (in_range) { cnt_evals += 1; cnt_resolved += r.candidates; cnt_scored += r.scored; }
        const uint32_t mykey = r.kept_total ? ((r.kept_total << QC_NODE_BITS) | key_low) : 0u;
        const uint32_t wkey = __reduce_max_sync(0xffffffffu, mykey);
        uint32_t* wk = wkeys + (p & 1u) * 32u;
        if (lane == 0) wk[warp] = wkey;
        __syncthreads();
        const uint32_t ckey = __reduce_max_sync(0xffffffffu, lane < n_warps ? wk[lane] : 0u);      // the CTA's best key, in every warp
        // slot p mod 4 of every CTA receives the arg-max of the cluster for workload p: each CTA's key arrives as a remote
        // atomic max that also completes 4 bytes of the receiver's mbarrier.  The slot of workload p + 2 is cleared here,
        // two exchanges before its first remote update can arrive and one after its last reader moved on.
        const uint32_t slot = ckeys_a + (p & 3u) * 4u, bar = bars + (p & 1u) * 8u;
        if (warp == 0) {
            if (lane == 0) {
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(ckeys_a + ((p + 2u) & 3u) * 4u), "r"(0u) : "memory");
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(C * 4u) : "memory");
            }
            if ((uint32_t)lane < C) qc_max_signal(qc_map_to_cta(slot, (uint32_t)lane), ckey, qc_map_to_cta(bar, (uint32_t)lane));
        }
        // While the keys travel: the lane that holds the CTA's best key is the only one of this CTA that can win, so it
        // works out the state its node would have after the commit and parks it; if it wins it only has to pick it up.
        const bool candidate = mykey != 0u && mykey == ckey;
        if (candidate) {
            LaneNode nx;
            qc_state(nx, st, T, allocated | ((r.w2 >> 16) & 0xffu), n_allocs + 1u);
            spec[0] = nx.avail; spec[1] = (uint32_t)nx.A; spec[2] = nx.rlo; spec[3] = nx.rhi;
            spec[4] = (uint32_t)nx.sum_free; spec[5] = (uint32_t)(nx.sum_free >> 32); spec[6] = (uint32_t)nx.util_bonus; spec[7] = (uint32_t)nx.balance;
            spec[8] = nx.arch_lo; spec[9] = nx.arch_hi; spec[10] = nx.avail_order; spec[11] = (uint32_t)nx.avail_prefix; spec[12] = (uint32_t)(nx.avail_prefix >> 32);
        }
        __syncwarp();
        bool arrived = true;
        if (lane == 0) arrived = qc_wait_phase(bar, (p >> 1) & 1u);
        arrived = __shfl_sync(0xffffffffu, arrived, 0);
        if (!arrived) { if (lane == 0 && counters) atomicAdd(&counters[3], 1ull); break; }      // (every warp of the cluster times out alike)
        uint32_t win;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(win) : "r"(slot) : "memory");
        if (PEERS) {
            // `win` is this rank's best.  Warp 0 of CTA 0 speaks for the rank: lane r sends the rank's word to rank r and waits
            // for rank r's; the outcome goes to every CTA of the cluster the way the keys did (st.async + mbarrier).
            const uint32_t gslot = ckeys_a + 64u + (p & 3u) * 4u, gbar = bars + 16u + (p & 1u) * 8u;
            if (tid == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(gbar), "r"(4u) : "memory");
            if (rank == 0 && tid < 32) {
                const uint32_t tag = (PQ.seq_base + p) % QC_SEQ_MOD + 1u;
                const uint32_t gnode = P.node_base + ((QC_MAX_NODES - 1u) - (win & (QC_MAX_NODES - 1u)));
                const unsigned long long own = win ? (((unsigned long long)(win >> QC_NODE_BITS) << 22) | (unsigned long long)(0x3FFFFFu - gnode)) : 0ull;
                unsigned long long best = own;
                bool ok = true;
                if ((uint32_t)lane < PQ.n_ranks && (uint32_t)lane != PQ.my_rank) {
                    const unsigned long long word = ((unsigned long long)tag << 40) | own;
                    unsigned long long* dst = PQ.recv[lane] + (PQ.my_rank * 4u + (tag & 3u));
                    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(word) : "memory");      // (the word is the whole message: nothing to release)
                    const unsigned long long* src = PQ.recv[PQ.my_rank] + ((uint32_t)lane * 4u + (tag & 3u));
                    unsigned long long v = 0;
                    const long long t0 = clock64();
                    for (;;) {
                        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(src) : "memory");
                        if ((uint32_t)(v >> 40) == tag) break;
                        if (clock64() - t0 > 20000000000ll) { ok = false; break; }      // ~10 s: a rank is gone
                    }
                    best = v & 0xFFFFFFFFFFull;
                }
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) { const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, off); best = o > best ? o : best; }
                ok = __all_sync(0xffffffffu, ok);
                const uint32_t gw = !ok ? QC_EXCHANGE_FAILED : best == 0ull ? 0u : best == own ? win : QC_PLACED_ELSEWHERE;
                if ((uint32_t)lane < C) qc_store_signal(qc_map_to_cta(gslot, (uint32_t)lane), gw, qc_map_to_cta(gbar, (uint32_t)lane));
            }
            __syncwarp();
            arrived = true;
            if (lane == 0) arrived = qc_wait_phase(gbar, (p >> 1) & 1u);
            arrived = __shfl_sync(0xffffffffu, arrived, 0);
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(win) : "r"(gslot) : "memory");
            if (!arrived || win == QC_EXCHANGE_FAILED) { if (lane == 0 && counters) atomicAdd(&counters[3], 1ull); break; }
        }
        if (win == 0u || (PEERS && win == QC_PLACED_ELSEWHERE)) {
            if (rank == 0 && tid == 0) {
                kgwe_placement pl;
                pl.node_idx = KGWE_NO_NODE; pl.total = 0; pl.topology = pl.resource = pl.balance = 0; pl.gpu_mask = 0; pl.flags = 0;
                pl.order = 0; pl.reserved = 0; pl.est_bandwidth_gbps = 0.0;
                out[p] = pl;
            }
            if (win == 0u) gang_failed = true;
        } else if (win == mykey) {
            // this lane owns the winner: tryScheduleOnNode + recordAllocation (scheduler.go:625-653, 695-707)
            const uint32_t mask = (r.w2 >> 16) & 0xffu;
            kgwe_placement pl;
            pl.node_idx = P.node_base + node; pl.total = (int32_t)r.w0; pl.topology = (int16_t)(r.w1 & 0xffffu); pl.resource = (int16_t)(r.w1 >> 16);
            pl.balance = (int16_t)(r.w2 & 0xffffu); pl.gpu_mask = (uint8_t)mask; pl.flags = (uint8_t)(r.w2 >> 24); pl.order = r.w3;
            pl.reserved = 0; pl.est_bandwidth_gbps = 0.0;      // EstimatedBandwidth is filled in after the queue (est_bandwidth_kernel)
            out[p] = pl;
            allocated |= mask; n_allocs += 1;
            gang_mask |= mask; gang_commits += 1;
            // the state parked above (tier, valid, integral never change)
            ln.avail = spec[0]; ln.A = (int)spec[1]; ln.rlo = spec[2]; ln.rhi = spec[3];
            ln.sum_free = (uint64_t)spec[4] | ((uint64_t)spec[5] << 32); ln.util_bonus = (int)spec[6]; ln.balance = (int)spec[7];
            ln.arch_lo = spec[8]; ln.arch_hi = spec[9]; ln.avail_order = spec[10]; ln.avail_prefix = (uint64_t)spec[11] | ((uint64_t)spec[12] << 32);
            memo_ok = 0; groups_ok = false;
            rec->allocated_mask = (uint8_t)allocated;      // the image in HBM follows
            rec->n_allocs = n_allocs;
        }
        if (gang_of && sgang[wi + 1] != sgang[wi]) {      // the gang's last member has been decided
            if (gang_failed && gang_commits) {             // take this node's commits for the gang back
                allocated &= ~gang_mask; n_allocs -= gang_commits;
                qc_state(ln, st, T, allocated, n_allocs);
                memo_ok = 0; groups_ok = false;
                rec->allocated_mask = (uint8_t)allocated;
                rec->n_allocs = n_allocs;
            }
            gang_failed = false; gang_mask = 0; gang_commits = 0;
        }
        __syncwarp();
    }
    cnt_scored = __reduce_add_sync(0xffffffffu, cnt_scored);
    cnt_resolved = __reduce_add_sync(0xffffffffu, cnt_resolved);
    cnt_evals = __reduce_add_sync(0xffffffffu, cnt_evals);
    if (lane == 0 && counters) {
        if (cnt_evals) atomicAdd(&counters[0], (unsigned long long)cnt_evals);
        if (cnt_scored) atomicAdd(&counters[1], (unsigned long long)cnt_scored);
        if (cnt_resolved) atomicAdd(&counters[2], (unsigned long long)cnt_resolved);
    }
    qc_cluster_arrive(); qc_cluster_wait();      // nobody leaves while a peer may still store into its shared memory; every decision is visible
    // EstimatedBandwidth of the decisions (scheduler.go:656-692): it depends only on link data that commits never change, so
    // it stays off the per-workload path and one CTA fills it in for the whole queue here.
    if (rank == 0)
        for (uint32_t p = (uint32_t)tid; p < n_pods; p += (uint32_t)T) {
            const uint32_t nd = out[p].node_idx;
            if (nd == KGWE_NO_NODE || nd - P.node_base >= n_nodes) continue;
            const uint32_t order = out[p].order;
            out[p].est_bandwidth_gbps = est_bandwidth(nodes + (nd - P.node_base), order, (int)((order >> 24) & 15));
        }
}

}  // namespace kgwe
