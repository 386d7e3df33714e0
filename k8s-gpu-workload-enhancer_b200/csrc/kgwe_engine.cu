// kgwe_engine.cu — host side of the engine and the C ABI of include/kgwe_engine.h.
//
// State (replaces TopologyAwareScheduler's fields, reference src/scheduler/scheduler.go:16-39):
//   * cluster image: TWO DevNode[max_nodes] images in HBM + a pinned host mirror.  Upserts and ledger calls
//     touch the mirror and two dirty lists (one per image); the next call that reads the image copies the
//     pending records into the SHADOW image on a copy stream, makes its own stream wait for that copy (an event,
//     no host synchronisation) and swaps the two (incremental and double-buffered, instead of
//     discovery.go:290-319's full rebuild): kernels still running on the old image are never written under;
//   * ledger: wl_id -> (node, gpu mask) on the host; its per-node projection
//     (allocated_mask, n_allocs) lives inside the node records, which is all the kernels need
//     (replaces the three full ledger scans per (pod,node) at scheduler.go:587-593, 562-566);
//   * scratch: arg-max key table, pod/hint/top-k staging, an optional score-matrix chunk.
// No CPU fallback: every scoring entry point launches the CUDA kernels or fails.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "kgwe_kernels.cuh"
#include "kgwe_mig.cuh"
#include "kgwe_queue.cuh"
#include "kgwe_pack.hpp"

using namespace kgwe;

namespace {

struct LedgerEntry { uint32_t node; uint8_t mask; };

template <typename T>
struct DevBuf {
    T* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t n) {
        if (n <= cap) return cudaSuccess;
        size_t want = std::max(n, cap * 2);
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
template <typename T>
struct PinBuf {
    T* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMallocHost(&p, n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

constexpr int kCounters = 4;      // device counters: [0] evaluations (K3), [1] candidates scored, [2] candidates by the reference's counting rule, [3] watchdog trips of the cluster queue kernel
constexpr size_t kScoreScratchBytes = 1ull << 30;   // internal score-matrix chunk (top-k > 1 without out_scores, host copies)
// Testing knob: KGWE_SCORE_SCRATCH_BYTES shrinks the chunk so the multi-pass path of kgwe_score_batch can be exercised on small inputs.
size_t score_scratch_bytes() {
    static const size_t v = [] { const char* s = getenv("KGWE_SCORE_SCRATCH_BYTES"); const long long x = s ? atoll(s) : 0; return x > 0 ? (size_t)x : kScoreScratchBytes; }();
    return v;
}

}  // namespace

struct kgwe_engine {
    std::mutex mu;
    kgwe_config cfg;
    Params params;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int sm_count = 148;
    int sched_blocks_per_sm = 1;
    int sched_tile_blocks_per_sm = 0;
    int qc_max_ctas = 0;              // largest cluster the cluster form of the queue kernel can be launched with here (0: none)
    long long max_total = 0;          // bound on |TotalScore| from the weights
    unsigned long long watchdog_seen = 0;     // h_counters[3] at the end of the previous queue call
    int k1_blocks_per_sm[4] = {5, 3, 2, 1};   // co-resident CTAs of K1's lean / GENERAL (int32 columns) / GENERAL (float64 columns) / TABLES variants

    DevNode* d_img[2] = {nullptr, nullptr};   // the two cluster images in HBM
    int front = 0;                    // the image scoring kernels read; the other one (shadow) receives the next updates
    DevNode* d_nodes = nullptr;       // == d_img[front]
    DevNode* h_nodes = nullptr;       // pinned mirror (device layout), authoritative
    std::vector<uint8_t> dirty_flag[2];       // per image: records that differ from the mirror
    std::vector<uint32_t> dirty[2];
    cudaStream_t cstream = nullptr;   // copy stream of the image updates
    cudaStream_t qstream = nullptr;   // counter read-back of kgwe_get_stats
    cudaEvent_t ev_copy = nullptr;    // the latest batch of image copies
    bool copy_pending = false;        // ev_copy may not have completed: the mirror records it reads must not change yet
    cudaEvent_t ev_use[2] = {nullptr, nullptr};   // after the latest launch that reads image i
    cudaEvent_t ev_last = nullptr;    // after the latest work of any entry point ...
    cudaStream_t last_stream = nullptr;       // ... on this stream (the engine's scratch is shared between calls)
    uint64_t epoch = 0;               // published image versions
    uint64_t host_evaluations = 0;    // evaluations counted on the host (K1: pods x present records)
    uint32_t n_nodes = 0;             // high-water mark of present records
    uint32_t n_present = 0;
    uint32_t n_nonuniform = 0;        // present records whose active links are not all equal (K1 then needs its GENERAL variant)
    uint32_t n_fractional = 0;        // ... of which with a fractional bandwidth (K1 then needs float64 matrix columns)
    std::unordered_map<uint64_t, LedgerEntry> ledger;

    DevBuf<kgwe_pod_req> d_pods;
    DevBuf<uint8_t> d_hint;
    DevBuf<unsigned long long> d_best;      // arg-max keys, kept all-zero between calls
    DevBuf<unsigned int> d_arrive;          // per-pod arrival counters of the queue kernel, kept all-zero between calls
    DevBuf<uint32_t> d_gang;                // gang numbers of the queue's pods (E5)
    DevBuf<kgwe_placement> d_topk;
    DevBuf<kgwe_node_score> d_scores;
    unsigned long long* d_counters = nullptr;   // kCounters words, see above
    PinBuf<kgwe_pod_req> h_pods;
    PinBuf<kgwe_placement> h_topk;
    PinBuf<uint8_t> h_hint;
    unsigned long long* h_counters = nullptr;
    unsigned int* d_done = nullptr;           // [0] latency path: CTA completion ticket; [1] the same for the fused P2P finalisation
    unsigned int* h_flag = nullptr;           // latency path: pinned word the kernel sets to flag_seq when the placements are written
    unsigned int flag_seq = 0;

    // K1 timing: event pairs recorded around each score_kernel launch, harvested by get_stats
    static constexpr int kTimers = 64;
    cudaEvent_t t0[kTimers] = {}, t1[kTimers] = {};
    int timers_pending = 0;

    kgwe_stats stats;
};

namespace {

// The message of the last failing call is kept per THREAD: kgwe_last_error hands out a pointer, and a message stored in
// the engine could be replaced (and freed) by another thread's failing call while the first thread still reads it.
thread_local std::string tl_last_error;

int fail(kgwe_engine* e, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    (void)e;
    tl_last_error = buf;
    return code;
}
#define CU(e, call) do { cudaError_t _c = (call); if (_c != cudaSuccess) return fail(e, KGWE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(_c)); } while (0)

int to_device_record(kgwe_engine* e, const kgwe_node_image& in, DevNode& out) {
    char msg[160];
    if (kgwe::to_device_record(in, out, msg, sizeof msg) != 0) return fail(e, KGWE_ERR_INVALID_ARG, "%s", msg);
    return KGWE_OK;
}

inline uint32_t rec_nonuniform(const DevNode& r) { return ((r.flags & NODE_PRESENT) && r.pad1[0] == 0) ? 1u : 0u; }
inline uint32_t rec_fractional(const DevNode& r) { return (rec_nonuniform(r) && !(r.flags & NODE_INTEGRAL)) ? 1u : 0u; }

void mark_dirty_image(kgwe_engine* e, int img, uint32_t idx) {
    if (!e->dirty_flag[img][idx]) { e->dirty_flag[img][idx] = 1; e->dirty[img].push_back(idx); }
}
// A record of the mirror changed: both images are stale there.
void mark_dirty(kgwe_engine* e, uint32_t idx) { mark_dirty_image(e, 0, idx); mark_dirty_image(e, 1, idx); }
// A queue kernel committed on the front image itself and the mirror has followed: only the shadow is stale.
void mark_dirty_shadow(kgwe_engine* e, uint32_t idx) { mark_dirty_image(e, 1 - e->front, idx); }

// Called before the host changes a record of the pinned mirror: a copy batch may still be reading it.
void wait_copy(kgwe_engine* e) {
    if (e->copy_pending) { cudaEventSynchronize(e->ev_copy); e->copy_pending = false; }
}

// Start of every entry point that reads the image on stream s:
//   - order s after the engine's previous work if that ran on another stream (scratch buffers are shared);
//   - if the front image is stale, bring the SHADOW image up to date on the copy stream (waiting, on the device, for
//     kernels that may still read it from the time it was front), let s wait for the copies, and swap.
// No host synchronisation: the caller's stream only gets event dependencies.
int publish(kgwe_engine* e, cudaStream_t s) {
    if (e->last_stream && e->last_stream != s) CU(e, cudaStreamWaitEvent(s, e->ev_last, 0));
    if (e->dirty[e->front].empty()) return KGWE_OK;
    const int sh = 1 - e->front;
    std::vector<uint32_t>& d = e->dirty[sh];
    CU(e, cudaStreamWaitEvent(e->cstream, e->ev_use[sh], 0));
    std::sort(d.begin(), d.end());
    size_t i = 0;
    while (i < d.size()) {
        size_t j = i + 1;
        while (j < d.size() && d[j] == d[j - 1] + 1) ++j;
        const uint32_t first = d[i], count = (uint32_t)(j - i);
        CU(e, cudaMemcpyAsync(e->d_img[sh] + first, e->h_nodes + first, (size_t)count * sizeof(DevNode), cudaMemcpyHostToDevice, e->cstream));
        e->stats.h2d_bytes += (uint64_t)count * sizeof(DevNode);
        i = j;
    }
    for (uint32_t idx : d) e->dirty_flag[sh][idx] = 0;
    d.clear();
    CU(e, cudaEventRecord(e->ev_copy, e->cstream));
    e->copy_pending = true;
    CU(e, cudaStreamWaitEvent(s, e->ev_copy, 0));
    e->front = sh; e->d_nodes = e->d_img[sh]; e->epoch++;
    return KGWE_OK;
}
int flush_dirty(kgwe_engine* e, cudaStream_t s) { return publish(e, s); }

// End of every entry point that launched work on s.
int note_use(kgwe_engine* e, cudaStream_t s) {
    CU(e, cudaEventRecord(e->ev_use[e->front], s));
    CU(e, cudaEventRecord(e->ev_last, s));
    e->last_stream = s;
    return KGWE_OK;
}

int validate_pods(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n) {
    for (uint32_t i = 0; i < n; ++i) {
        if (pods[i].gpu_count < 0) return fail(e, KGWE_ERR_INVALID_ARG, "pod %u: gpu_count %d < 0", i, pods[i].gpu_count);
        if (pods[i].min_memory_gb < 0 || pods[i].min_memory_gb > (1 << 20)) return fail(e, KGWE_ERR_INVALID_ARG, "pod %u: min_memory_gb %d outside [0, 2^20]", i, pods[i].min_memory_gb);
        if (!(pods[i].min_bandwidth_gbps >= 0.0)) return fail(e, KGWE_ERR_INVALID_ARG, "pod %u: min_bandwidth_gbps must be >= 0", i);
    }
    return KGWE_OK;
}

uint32_t hint_row_bytes(const kgwe_engine* e) { return (e->n_nodes + 7) / 8; }

// Latency histogram of kgwe_stats: `n` observations of `us` microseconds each.
void observe_latency(kgwe_engine* e, double us, uint64_t n) {
    static const double edges[KGWE_LATENCY_BUCKETS - 1] = {10, 25, 50, 100, 250, 500, 1000, 2500, 5000};
    int b = 0;
    while (b < KGWE_LATENCY_BUCKETS - 1 && us > edges[b]) ++b;
    e->stats.latency_us_bucket[b] += n; e->stats.latency_count += n; e->stats.latency_sum_us += us * (double)n;
}
struct CallTimer {
    kgwe_engine* e; uint64_t n; std::chrono::steady_clock::time_point t0;
    CallTimer(kgwe_engine* e_, uint64_t n_) : e(e_), n(n_), t0(std::chrono::steady_clock::now()) {}
    void done(bool per_item) {
        const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
        if (n) observe_latency(e, per_item ? us / (double)n : us, per_item ? n : 1);
    }
};
size_t sched_tile_smem_bytes() { return (size_t)TILE_NODES * (sizeof(DevNode) + sizeof(NodeFast) + sizeof(NodeCtx) + 64) + 16; }

// ---- K1 launch ------------------------------------------------------------------------------------
// Kernel variant: the lean one when every present record is uniform (all active links of a node equal - any
// NVSwitch node, any single-generation mesh), GENERAL as soon as one record is not (matrix columns in shared
// memory), TABLES when exhaustive-mode pods may be in the batch (subset-key columns; implies GENERAL).
struct K1Plan { bool general, tables; uint32_t mode, chunks, ppc, n_tiles, mat_units, groups; size_t smem; };

uint32_t k1_mode(const kgwe_engine* e) {
    return (e->cfg.flags & KGWE_CFG_MEMOISE_GROUPS) ? K1_MODE_MEMO : (e->cfg.flags & KGWE_CFG_PRUNE_TIED_STARTS) ? K1_MODE_PRUNE : K1_MODE_FULL;
}

// Pod chunks: whole waves of co-resident CTAs.  Every CTA pays its set-up (node state, pod decode) once, so longer
// chunks amortise it better, but CTAs differ in cost (absent or full nodes finish early) and several shorter waves even
// that out: measured on the headline workload, chunks of about 270 pods are the best trade at every shard size
// (1.30 -> 1.24 ms at 10,000 nodes with four waves; a 1,250-node shard keeps its single wave of 136-pod chunks).
K1Plan plan_k1(const kgwe_engine* e, uint32_t n_pods, bool tables, bool fused_tail = false) {
    K1Plan pl;
    pl.tables = tables || (e->cfg.flags & KGWE_CFG_SUBSET_TABLES);
    pl.general = pl.tables || e->n_nonuniform > 0;
    pl.mat_units = !pl.general ? 0u : e->n_fractional > 0 ? 32u : 16u;      // float64 rows only when some non-uniform record has fractional bandwidths
    pl.mode = k1_mode(e);
    // the variants with big per-node tables run two warp groups per CTA that share them (kgwe_score.cuh); not on the latency path,
    // whose in-kernel finalisation is written for one group
    pl.groups = ((pl.tables || pl.mat_units == 32u) && !fused_tail) ? (pl.tables ? 4u : (uint32_t)K1_MAX_GROUPS) : 1u;
    pl.n_tiles = (e->n_nodes + K1_TILE - 1) / K1_TILE;
    const uint32_t max_ppc = pl.tables ? 384u : (uint32_t)MAX_PODS_PER_CHUNK;      // the subset columns leave little room for the pods' tables
    const uint32_t slots = (uint32_t)e->sm_count * (uint32_t)e->k1_blocks_per_sm[pl.tables ? 3 : !pl.general ? 0 : pl.mat_units == 16 ? 1 : 2];
    uint32_t chunks = 1, ppc = n_pods;
    uint32_t first_waves = 1;
    if (!pl.tables) {      // (the subset-table variant has the heaviest set-up: it keeps the longest chunks it can hold)
        const uint32_t one_wave_chunks = std::max(1u, std::min(n_pods, slots / pl.n_tiles));
        const uint32_t one_wave_ppc = (n_pods + one_wave_chunks - 1) / one_wave_chunks;
        first_waves = std::max(1u, (one_wave_ppc + 135u * pl.groups) / (270u * pl.groups));
    }
    for (uint32_t waves = first_waves; waves <= 256; ++waves) {
        chunks = std::max(1u, std::min(n_pods, slots * waves / pl.n_tiles));
        ppc = (n_pods + chunks - 1) / chunks;
        if (ppc <= max_ppc) break;      // (one wave is fine: all CTAs of a launch do the same amount of work)
    }
    if (ppc > max_ppc) ppc = max_ppc;
    pl.ppc = ppc;
    pl.chunks = (n_pods + ppc - 1) / ppc;
    pl.smem = k1_layout(ppc, pl.mat_units, pl.tables, pl.mode == K1_MODE_MEMO, pl.groups).total;
    return pl;
}

// Launch K1 alone: scores into d_scores (nullable), arg-max keys into `best` (nullable).
int launch_k1(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods, const uint8_t* d_hint, kgwe_node_score* d_scores,
              unsigned long long* best, cudaStream_t s, bool tables, const FusedTail& tail, bool timed) {
    const K1Plan pl = plan_k1(e, n_pods, tables, tail.out != nullptr);
    if (pl.chunks > 65535u) return fail(e, KGWE_ERR_TOO_LARGE, "too many pods in one launch (%u)", n_pods);
    const dim3 grid(pl.n_tiles, pl.chunks);
    timed = timed && e->timers_pending < kgwe_engine::kTimers;
    if (timed) CU(e, cudaEventRecord(e->t0[e->timers_pending], s));
#define KGWE_LAUNCH_SCORE(G, T) score_kernel<G, T><<<grid, K1_THREADS * pl.groups, pl.smem, s>>>(e->d_nodes, e->n_nodes, d_pods, n_pods, pl.ppc, d_hint, hint_row_bytes(e), \
                                                                                    d_scores, best, e->params, e->d_counters, pl.mode, pl.mat_units, tail)
    if (pl.tables) KGWE_LAUNCH_SCORE(true, true);
    else if (pl.general) KGWE_LAUNCH_SCORE(true, false);
    else KGWE_LAUNCH_SCORE(false, false);
#undef KGWE_LAUNCH_SCORE
    CU(e, cudaGetLastError());
    if (timed) { CU(e, cudaEventRecord(e->t1[e->timers_pending], s)); e->timers_pending++; }
    e->stats.kernel_launches++;
    e->host_evaluations += (uint64_t)n_pods * e->n_present;
    e->stats.evaluations += (uint64_t)n_pods * e->n_present;
    return KGWE_OK;
}

// Arg-max key table of the engine for n_pods pods (grown on demand, kept all-zero between calls).
int ensure_best(kgwe_engine* e, uint32_t n_pods, cudaStream_t s) {
    const size_t had = e->d_best.cap;
    CU(e, e->d_best.ensure(n_pods));
    if (e->d_best.cap != had) CU(e, cudaMemsetAsync(e->d_best.p, 0, e->d_best.cap * sizeof(unsigned long long), s));
    return KGWE_OK;
}

// Launch K1 (+K1b / K2) for n_pods pods whose descriptors, hint rows and outputs are in
// device memory.  d_scores may be null unless topk > 1.
int launch_score(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods, const uint8_t* d_hint,
                 kgwe_node_score* d_scores, uint32_t topk, kgwe_placement* d_topk, cudaStream_t s, bool tables = false,
                 bool fused_tail = false) {
    FusedTail tail{};
    if (fused_tail) {
        tail.out = d_topk; tail.done = e->d_done; tail.flag = e->h_flag; tail.seq = ++e->flag_seq;
        if (n_pods <= 4) { tail.n_inline = n_pods; memcpy(tail.inline_pods, d_pods, (size_t)n_pods * sizeof(kgwe_pod_req)); }   // d_pods is the pinned HOST copy on this path
    }
    if (n_pods == 0) return KGWE_OK;
    if (e->n_nodes == 0) {
        // empty image (host entry points answer this themselves): every placement is "no node"
        if (topk && d_topk && !fused_tail) {
            const size_t n = (size_t)n_pods * topk;
            no_node_fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d_topk, n);
            CU(e, cudaGetLastError());
            e->stats.kernel_launches++;
        }
        return KGWE_OK;
    }
    unsigned long long* best = nullptr;
    if (topk == 1) {
        int rc = ensure_best(e, n_pods, s);
        if (rc != KGWE_OK) return rc;
        best = e->d_best.p;
    }
    int rc = launch_k1(e, d_pods, n_pods, d_hint, d_scores, best, s, tables, tail, !fused_tail);
    if (rc != KGWE_OK) return rc;
    if (fused_tail) return KGWE_OK;
    if (topk == 1) {
        const uint32_t threads = 256, pods_per_block = threads / 8;
        finalize_top1_kernel<<<(n_pods + pods_per_block - 1) / pods_per_block, threads, 0, s>>>(
            e->d_nodes, e->n_nodes, d_pods, n_pods, d_hint, hint_row_bytes(e), d_scores, best, d_topk, 1, e->params);
        CU(e, cudaGetLastError());
        e->stats.kernel_launches++;
    } else if (topk > 1) {
        topk_select_kernel<<<n_pods, 256, 0, s>>>(e->d_nodes, e->n_nodes, d_scores, topk, d_topk, e->params);
        CU(e, cudaGetLastError());
        e->stats.kernel_launches++;
    }
    return KGWE_OK;
}

// Device counters are monotonic (never reset): reading them needs no exclusion against running kernels.
int pull_counters(kgwe_engine* e, cudaStream_t s) {
    CU(e, cudaMemcpyAsync(e->h_counters, e->d_counters, kCounters * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    return KGWE_OK;
}
void fold_counters(kgwe_engine* e) {      // after the copy of pull_counters has completed
    e->stats.evaluations = e->host_evaluations + e->h_counters[0];
    e->stats.candidates = e->h_counters[1];
    e->stats.candidates_resolved = e->h_counters[2];
}

int ledger_apply(kgwe_engine* e, uint64_t wl_id, uint32_t idx, uint8_t mask, bool touch_device) {
    // recordAllocation: the same UID overwrites its previous entry (scheduler.go:699)
    auto it = e->ledger.find(wl_id);
    if (it != e->ledger.end()) {
        const uint32_t old_node = it->second.node;
        e->ledger.erase(it);
        uint8_t m = 0; uint32_t n = 0;
        for (auto& kv : e->ledger) if (kv.second.node == old_node) { m |= kv.second.mask; n++; }
        e->h_nodes[old_node].allocated_mask = m; e->h_nodes[old_node].n_allocs = n;
        mark_dirty(e, old_node);
    }
    DevNode& rec = e->h_nodes[idx];
    rec.allocated_mask |= mask; rec.n_allocs++;
    e->ledger[wl_id] = LedgerEntry{idx, mask};
    if (touch_device) mark_dirty(e, idx);
    return KGWE_OK;
}

}  // namespace

extern "C" {

#define KGWE_STR2(x) #x
#define KGWE_STR(x) KGWE_STR2(x)
const char* kgwe_version(void) { return "kgwe-b200 0.2 (sm_100a) abi " KGWE_STR(KGWE_ABI_VERSION); }

void kgwe_config_default(kgwe_config* c) {
    if (!c) return;
    memset(c, 0, sizeof *c);
    c->struct_size = sizeof *c; c->device = 0; c->max_nodes = 16384; c->node_base = 0;
    c->topology_weight = 40; c->resource_weight = 35; c->balance_weight = 25; c->hint_bonus = 10;
    c->nvlink_norm_gbps = 900.0;
}

int kgwe_engine_create(const kgwe_config* cfg, kgwe_engine** out) {
    if (!cfg || !out || cfg->struct_size != sizeof(kgwe_config) || cfg->max_nodes == 0) return KGWE_ERR_INVALID_ARG;
    auto bad_w = [](int w) { return w < -1000 || w > 1000; };
    if (bad_w(cfg->topology_weight) || bad_w(cfg->resource_weight) || bad_w(cfg->balance_weight) || bad_w(cfg->hint_bonus)) return KGWE_ERR_INVALID_ARG;
    if (!(cfg->nvlink_norm_gbps >= 400.0) || !(cfg->nvlink_norm_gbps <= 1e6)) return KGWE_ERR_INVALID_ARG;   // keeps TopologyScore inside int16
    {   // K1 ranks with 32-bit keys (TotalScore << 7 | node-in-tile): with TopologyScore <= 12,550 (links <= 1e5 GB/s at norm >= 400),
        // Resource / Balance <= 100 and these weights, TotalScore stays below 2^24
        auto ab = [](long long v) { return v < 0 ? -v : v; };
        const long long max_total = (12550 * ab(cfg->topology_weight) + 100 * ab(cfg->resource_weight) + 100 * ab(cfg->balance_weight)) / 100 + ab(cfg->hint_bonus);
        if (max_total >= (1ll << 24)) return KGWE_ERR_INVALID_ARG;
    }
    *out = nullptr;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0 || cfg->device < 0 || cfg->device >= n_dev) return KGWE_ERR_NO_DEVICE;
    if (cudaSetDevice(cfg->device) != cudaSuccess) return KGWE_ERR_NO_DEVICE;
    kgwe_engine* e = new (std::nothrow) kgwe_engine();
    if (!e) return KGWE_ERR_NO_MEMORY;
    e->cfg = *cfg;
    e->params = Params{cfg->topology_weight, cfg->resource_weight, cfg->balance_weight, cfg->hint_bonus, cfg->nvlink_norm_gbps, cfg->node_base};
    memset(&e->stats, 0, sizeof e->stats);
    bool ok = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess
           && cudaEventCreate(&e->ev0) == cudaSuccess && cudaEventCreate(&e->ev1) == cudaSuccess
           && cudaMalloc(&e->d_img[0], (size_t)cfg->max_nodes * sizeof(DevNode)) == cudaSuccess && cudaMalloc(&e->d_img[1], (size_t)cfg->max_nodes * sizeof(DevNode)) == cudaSuccess
           && cudaStreamCreateWithFlags(&e->cstream, cudaStreamNonBlocking) == cudaSuccess && cudaStreamCreateWithFlags(&e->qstream, cudaStreamNonBlocking) == cudaSuccess
           && cudaEventCreateWithFlags(&e->ev_copy, cudaEventDisableTiming) == cudaSuccess && cudaEventCreateWithFlags(&e->ev_last, cudaEventDisableTiming) == cudaSuccess
           && cudaEventCreateWithFlags(&e->ev_use[0], cudaEventDisableTiming) == cudaSuccess && cudaEventCreateWithFlags(&e->ev_use[1], cudaEventDisableTiming) == cudaSuccess
           && cudaMallocHost(&e->h_nodes, (size_t)cfg->max_nodes * sizeof(DevNode)) == cudaSuccess
           && cudaMalloc(&e->d_counters, kCounters * sizeof(unsigned long long)) == cudaSuccess
           && cudaMallocHost(&e->h_counters, kCounters * sizeof(unsigned long long)) == cudaSuccess
           && cudaMalloc(&e->d_done, 2 * sizeof(unsigned int)) == cudaSuccess && cudaMemset(e->d_done, 0, 2 * sizeof(unsigned int)) == cudaSuccess
           && cudaMallocHost(&e->h_flag, sizeof(unsigned int)) == cudaSuccess
           && cudaMemset(e->d_img[0], 0, (size_t)cfg->max_nodes * sizeof(DevNode)) == cudaSuccess && cudaMemset(e->d_img[1], 0, (size_t)cfg->max_nodes * sizeof(DevNode)) == cudaSuccess
           && cudaMemset(e->d_counters, 0, kCounters * sizeof(unsigned long long)) == cudaSuccess;
    if (ok) {
        memset(e->h_nodes, 0, (size_t)cfg->max_nodes * sizeof(DevNode));
        for (int i = 0; i < kCounters; ++i) e->h_counters[i] = 0;
        *e->h_flag = 0;
        e->dirty_flag[0].assign(cfg->max_nodes, 0); e->dirty_flag[1].assign(cfg->max_nodes, 0);
        e->d_nodes = e->d_img[0];
        cudaDeviceGetAttribute(&e->sm_count, cudaDevAttrMultiProcessorCount, cfg->device);
        for (int i = 0; i < kgwe_engine::kTimers && ok; ++i) ok = cudaEventCreate(&e->t0[i]) == cudaSuccess && cudaEventCreate(&e->t1[i]) == cudaSuccess;
        {
            const int lean = (int)k1_layout(MAX_PODS_PER_CHUNK, 0, false, true).total, gen = (int)k1_layout(MAX_PODS_PER_CHUNK, 32, false, true, K1_MAX_GROUPS).total,
                      tab = (int)k1_layout(384, 32, true, true, 4).total;
            ok = ok && cudaFuncSetAttribute(score_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lean) == cudaSuccess
                    && cudaFuncSetAttribute(score_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, gen) == cudaSuccess
                    && cudaFuncSetAttribute(score_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tab) == cudaSuccess;
            // co-resident CTAs at a typical chunk size (plan_k1 sizes its waves with these)
            int b = 0;
            if (ok && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, score_kernel<false, false>, K1_THREADS, k1_layout(512, 0, false, false).total) == cudaSuccess && b > 0) e->k1_blocks_per_sm[0] = b;
            if (ok && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, score_kernel<true, false>, K1_THREADS, k1_layout(512, 16, false, false).total) == cudaSuccess && b > 0) e->k1_blocks_per_sm[1] = b;
            if (ok && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, score_kernel<true, false>, K1_THREADS * K1_MAX_GROUPS, k1_layout(540, 32, false, false, K1_MAX_GROUPS).total) == cudaSuccess && b > 0) e->k1_blocks_per_sm[2] = b;
            if (ok && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, score_kernel<true, true>, K1_THREADS * 4, k1_layout(384, 32, true, false, 4).total) == cudaSuccess && b > 0) e->k1_blocks_per_sm[3] = b;
        }
        int bps = 0;
        if (ok && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, schedule_kernel, SCORE_THREADS, 0) == cudaSuccess && bps > 0) e->sched_blocks_per_sm = bps;
        ok = ok && cudaFuncSetAttribute(schedule_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sched_tile_smem_bytes()) == cudaSuccess;
        bps = 0;
        if (ok && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, schedule_tile_kernel, SCORE_THREADS, sched_tile_smem_bytes()) == cudaSuccess && bps > 0) e->sched_tile_blocks_per_sm = bps;
    }
    if (ok) {   // the cluster form of the queue kernel (kgwe_queue.cuh): the largest cluster this device schedules, 16 CTAs if it can
        auto ab = [](long long v) { return v < 0 ? -v : v; };
        e->max_total = (12550 * ab(cfg->topology_weight) + 100 * ab(cfg->resource_weight) + 100 * ab(cfg->balance_weight)) / 100 + ab(cfg->hint_bonus);
        const int smem = (int)qc_layout(QC_MAX_THREADS).total;
        const char* form = getenv("KGWE_QUEUE_FORM");      // "tile" keeps the grid-wide kernels (A/B runs, tests)
        std::vector<uint64_t> pf(256); std::vector<uint32_t> od(256);
        qc_first_tables(pf.data(), od.data());
        if (!(form && !strcmp(form, "tile")) && cudaFuncSetAttribute(schedule_cluster_kernel<QC_MAX_THREADS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) == cudaSuccess
            && cudaFuncSetAttribute(schedule_cluster_kernel<QC_MAX_THREADS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) == cudaSuccess
            && cudaFuncSetAttribute(schedule_cluster_kernel<QC_SMALL_THREADS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qc_layout(QC_SMALL_THREADS).total) == cudaSuccess
            && cudaFuncSetAttribute(schedule_cluster_kernel<QC_SMALL_THREADS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qc_layout(QC_SMALL_THREADS).total) == cudaSuccess
            && cudaMemcpyToSymbol(qc_prefix_lut, pf.data(), 256 * sizeof(uint64_t)) == cudaSuccess && cudaMemcpyToSymbol(qc_order_lut, od.data(), 256 * sizeof(uint32_t)) == cudaSuccess) {
            cudaFuncSetAttribute(schedule_cluster_kernel<QC_MAX_THREADS, false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            cudaFuncSetAttribute(schedule_cluster_kernel<QC_SMALL_THREADS, false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            cudaFuncSetAttribute(schedule_cluster_kernel<QC_MAX_THREADS, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            cudaFuncSetAttribute(schedule_cluster_kernel<QC_SMALL_THREADS, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            for (int c = QC_MAX_CTAS; c >= 1 && !e->qc_max_ctas; c >>= 1) {
                cudaLaunchConfig_t lc = {};
                lc.gridDim = dim3(c); lc.blockDim = dim3(QC_MAX_THREADS); lc.dynamicSmemBytes = smem;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = c; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                lc.attrs = at; lc.numAttrs = 1;
                int n = 0;
                if (cudaOccupancyMaxActiveClusters(&n, schedule_cluster_kernel<QC_MAX_THREADS, true>, &lc) == cudaSuccess && n >= 1) e->qc_max_ctas = c;
            }
        }
        cudaGetLastError();
    }
    if (!ok) { cudaGetLastError(); kgwe_engine_destroy(e); return KGWE_ERR_CUDA; }
    *out = e;
    return KGWE_OK;
}

int kgwe_engine_destroy(kgwe_engine* e) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    cudaSetDevice(e->cfg.device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    e->d_pods.release(); e->d_hint.release(); e->d_best.release(); e->d_arrive.release(); e->d_gang.release(); e->d_topk.release(); e->d_scores.release();
    e->h_pods.release(); e->h_topk.release(); e->h_hint.release();
    if (e->cstream) cudaStreamSynchronize(e->cstream);
    if (e->qstream) cudaStreamSynchronize(e->qstream);
    for (int i = 0; i < 2; ++i) { if (e->d_img[i]) cudaFree(e->d_img[i]); if (e->ev_use[i]) cudaEventDestroy(e->ev_use[i]); }
    if (e->ev_copy) cudaEventDestroy(e->ev_copy);
    if (e->ev_last) cudaEventDestroy(e->ev_last);
    if (e->cstream) cudaStreamDestroy(e->cstream);
    if (e->qstream) cudaStreamDestroy(e->qstream);
    if (e->h_nodes) cudaFreeHost(e->h_nodes);
    if (e->d_counters) cudaFree(e->d_counters);
    if (e->h_counters) cudaFreeHost(e->h_counters);
    if (e->d_done) cudaFree(e->d_done);
    if (e->h_flag) cudaFreeHost(e->h_flag);
    for (int i = 0; i < kgwe_engine::kTimers; ++i) { if (e->t0[i]) cudaEventDestroy(e->t0[i]); if (e->t1[i]) cudaEventDestroy(e->t1[i]); }
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
    return KGWE_OK;
}

// Ledger projection of one node from the ledger itself.
static void recount_node(kgwe_engine* e, uint32_t node) {
    uint8_t m = 0; uint32_t n = 0;
    for (auto& kv : e->ledger) if (kv.second.node == node) { m |= kv.second.mask; n++; }
    e->h_nodes[node].allocated_mask = m; e->h_nodes[node].n_allocs = n;
}

// One record into the mirror (mutex held).  old_to_new (nullable): the GPU list of the node changed - see kgwe_node_upsert_remap.
static int upsert_one(kgwe_engine* e, uint32_t idx, const kgwe_node_image& img, const int8_t* old_to_new) {
    DevNode rec;
    int rc = to_device_record(e, img, rec);
    if (rc != KGWE_OK) return rc;
    if (old_to_new)
        for (int g = 0; g < KGWE_MAX_GPUS; ++g)
            if (old_to_new[g] >= (int)rec.n_gpus) return fail(e, KGWE_ERR_INVALID_ARG, "old_to_new[%d] = %d names a GPU the new record does not have (%u GPUs)", g, old_to_new[g], rec.n_gpus);
    wait_copy(e);
    DevNode& dst = e->h_nodes[idx];
    const bool was = dst.flags & 1;
    if (old_to_new) {
        // the ledger names GPUs by their index in the OLD list: move every entry of this node to the new indices
        // (the reference keys allocations by GPU UUID, scheduler.go:586-598, so they follow the GPU, not the slot)
        for (auto& kv : e->ledger) {
            if (kv.second.node != idx) continue;
            uint8_t m = 0;
            for (int g = 0; g < KGWE_MAX_GPUS; ++g)
                if ((kv.second.mask >> g & 1) && old_to_new[g] >= 0) m |= (uint8_t)(1u << old_to_new[g]);
            kv.second.mask = m;              // an entry whose GPUs all vanished still counts as a workload on the node (scheduler.go:561-566)
        }
    }
    e->n_nonuniform += rec_nonuniform(rec) - rec_nonuniform(dst);
    e->n_fractional += rec_fractional(rec) - rec_fractional(dst);
    if (was && !old_to_new) { rec.allocated_mask = dst.allocated_mask; rec.n_allocs = dst.n_allocs; dst = rec; }     // ledger-owned fields survive a refresh
    else {
        // a node (re)appearing under an index the ledger still names gets its allocations back, as the
        // reference's name-keyed ledger would match them again (scheduler.go:587-593)
        dst = rec;
        recount_node(e, idx);
    }
    if (!was) e->n_present++;
    mark_dirty(e, idx);
    e->n_nodes = std::max(e->n_nodes, idx + 1);
    return KGWE_OK;
}

int kgwe_nodes_upsert(kgwe_engine* e, uint32_t first, uint32_t count, const kgwe_node_image* imgs) {
    if (!e || (!imgs && count)) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if ((uint64_t)first + count > e->cfg.max_nodes) return fail(e, KGWE_ERR_OUT_OF_RANGE, "nodes [%u,%u) exceed max_nodes %u", first, first + count, e->cfg.max_nodes);
    for (uint32_t i = 0; i < count; ++i) {
        int rc = upsert_one(e, first + i, imgs[i], nullptr);
        if (rc != KGWE_OK) return rc;
    }
    return KGWE_OK;
}

int kgwe_node_upsert(kgwe_engine* e, uint32_t idx, const kgwe_node_image* img) { return kgwe_nodes_upsert(e, idx, 1, img); }

int kgwe_node_upsert_remap(kgwe_engine* e, uint32_t idx, const kgwe_node_image* img, const int8_t* old_to_new) {
    if (!e || !img || !old_to_new) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (idx >= e->cfg.max_nodes) return fail(e, KGWE_ERR_OUT_OF_RANGE, "node %u out of range", idx);
    uint32_t seen = 0;
    for (int g = 0; g < KGWE_MAX_GPUS; ++g) {
        if (old_to_new[g] < 0) continue;
        if (old_to_new[g] >= KGWE_MAX_GPUS || (seen >> old_to_new[g] & 1)) return fail(e, KGWE_ERR_INVALID_ARG, "old_to_new is not an injective map into 0..%d", KGWE_MAX_GPUS - 1);
        seen |= 1u << old_to_new[g];
    }
    return upsert_one(e, idx, *img, old_to_new);
}

int kgwe_node_remove(kgwe_engine* e, uint32_t idx) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (idx >= e->cfg.max_nodes) return fail(e, KGWE_ERR_OUT_OF_RANGE, "node %u out of range", idx);
    DevNode& dst = e->h_nodes[idx];
    if (!(dst.flags & 1)) return fail(e, KGWE_ERR_NOT_FOUND, "node %u not present", idx);
    // the reference deletes the map entry (discovery.go:597-600); ledger entries naming the node stay
    // in the ledger (they are keyed by node name there) and simply stop matching any node - until a
    // node comes back under the same index (see kgwe_nodes_upsert).  n_nodes is a high-water mark and
    // does not shrink, so the stride of score matrices and hint rows stays what the host knows.
    wait_copy(e);
    e->n_nonuniform -= rec_nonuniform(dst);
    e->n_fractional -= rec_fractional(dst);
    memset(&dst, 0, sizeof dst);
    e->n_present--;
    mark_dirty(e, idx);
    return KGWE_OK;
}

int kgwe_alloc_commit(kgwe_engine* e, uint64_t wl_id, uint32_t idx, uint8_t gpu_mask) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (idx >= e->cfg.max_nodes) return fail(e, KGWE_ERR_OUT_OF_RANGE, "node %u out of range", idx);
    const DevNode& rec = e->h_nodes[idx];
    if (rec.flags & NODE_PRESENT) {
        // tryScheduleOnNode re-checks that none of the selected GPUs is in any allocation (scheduler.go:634-640); the entry
        // this workload id already holds does not count, recordAllocation replaces it (:699)
        if (gpu_mask >> rec.n_gpus) return fail(e, KGWE_ERR_INVALID_ARG, "gpu_mask 0x%02x names a GPU node %u does not have (%u GPUs)", gpu_mask, idx, rec.n_gpus);
        uint8_t others = 0;
        for (auto& kv : e->ledger) if (kv.second.node == idx && kv.first != wl_id) others |= kv.second.mask;
        if (others & gpu_mask) return fail(e, KGWE_ERR_CONFLICT, "GPUs 0x%02x of node %u are already allocated", others & gpu_mask, idx);
    }
    wait_copy(e);
    return ledger_apply(e, wl_id, idx, gpu_mask, true);
}

int kgwe_alloc_release(kgwe_engine* e, uint64_t wl_id) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    auto it = e->ledger.find(wl_id);
    if (it == e->ledger.end()) return fail(e, KGWE_ERR_NOT_FOUND, "allocation not found for workload %llu", (unsigned long long)wl_id);
    const uint32_t node = it->second.node;
    e->ledger.erase(it);
    wait_copy(e);
    recount_node(e, node);
    mark_dirty(e, node);
    return KGWE_OK;
}

int kgwe_score_batch_device(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods, const uint8_t* d_hint,
                            kgwe_node_score* d_out, uint32_t topk, kgwe_placement* d_topk, void* cuda_stream) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (topk > KGWE_MAX_TOPK || (topk && !d_topk) || (n_pods && !d_pods)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : e->stream;
    int rc = flush_dirty(e, s);
    if (rc != KGWE_OK) return rc;
    kgwe_node_score* scores = d_out;
    if (topk > 1 && !scores) {
        if ((size_t)n_pods * e->n_nodes * sizeof(kgwe_node_score) > 8 * kScoreScratchBytes) return fail(e, KGWE_ERR_TOO_LARGE, "top-k > 1 without d_out_scores needs a %zu-byte scratch matrix; pass d_out_scores or split the batch", (size_t)n_pods * e->n_nodes * sizeof(kgwe_node_score));
        CU(e, e->d_scores.ensure((size_t)n_pods * e->n_nodes));
        scores = e->d_scores.p;
    }
    rc = launch_score(e, d_pods, n_pods, d_hint, scores, topk, d_topk, s);
    if (rc != KGWE_OK) return rc;
    return note_use(e, s);
}

int kgwe_score_batch_device_p2p(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods, const uint8_t* d_hint,
                                kgwe_node_score* d_out, const uint64_t* peer_bufs, uint32_t n_ranks, uint32_t my_rank, void* cuda_stream) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (!peer_bufs || n_ranks == 0 || n_ranks > 8 || my_rank >= n_ranks || (n_pods && !d_pods)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments (1..8 ranks)");
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : e->stream;
    int rc = flush_dirty(e, s);
    if (rc != KGWE_OK) return rc;
    if (n_pods == 0) return KGWE_OK;
    if (e->n_nodes == 0) return fail(e, KGWE_ERR_INVALID_ARG, "the cluster image of this shard is empty");
    // K1 exactly as in launch_score(topk = 1), then the fused finalise + scatter
    rc = ensure_best(e, n_pods, s);
    if (rc != KGWE_OK) return rc;
    unsigned long long* best = e->d_best.p;
    rc = launch_k1(e, d_pods, n_pods, d_hint, d_out, best, s, false, FusedTail{}, true);
    if (rc != KGWE_OK) return rc;
    PeerBufs peers;
    for (uint32_t r = 0; r < 8; ++r) peers.p[r] = r < n_ranks ? reinterpret_cast<kgwe_placement*>(peer_bufs[r]) : nullptr;
    finalize_top1_p2p_kernel<<<(n_pods + 31) / 32, 256, 0, s>>>(e->d_nodes, e->n_nodes, d_pods, n_pods, d_hint, hint_row_bytes(e), d_out, best, peers, n_ranks, my_rank, e->params,
                                                                PeerFlags{}, 0u, nullptr);
    CU(e, cudaGetLastError());
    e->stats.kernel_launches++;
    return note_use(e, s);
}

int kgwe_score_merge_p2p(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods, const uint8_t* d_hint, kgwe_node_score* d_out,
                         const uint64_t* peer_bufs, const uint64_t* peer_flags, uint32_t n_ranks, uint32_t my_rank, uint32_t epoch,
                         kgwe_placement* d_merged, void* cuda_stream) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (!peer_bufs || !peer_flags || !d_merged || n_ranks < 2 || n_ranks > 8 || my_rank >= n_ranks || epoch == 0 || (n_pods && !d_pods))
        return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments (2..8 ranks, epoch >= 1)");
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : e->stream;
    int rc = publish(e, s);
    if (rc != KGWE_OK) return rc;
    if (n_pods == 0) return KGWE_OK;
    if (e->n_nodes == 0) return fail(e, KGWE_ERR_INVALID_ARG, "the cluster image of this shard is empty");
    rc = ensure_best(e, n_pods, s);
    if (rc != KGWE_OK) return rc;
    unsigned long long* best = e->d_best.p;
    rc = launch_k1(e, d_pods, n_pods, d_hint, d_out, best, s, false, FusedTail{}, true);
    if (rc != KGWE_OK) return rc;
    // double-buffered record slots: a rank one step ahead writes the other slot (its step after that cannot start
    // before it has seen this rank's flag of the step in between, i.e. after this rank's merge of the current one)
    const size_t slot_off = (size_t)(epoch & 1u) * n_ranks * n_pods;
    PeerBufs peers; PeerFlags flags;
    for (uint32_t r = 0; r < 8; ++r) {
        peers.p[r] = r < n_ranks ? reinterpret_cast<kgwe_placement*>(peer_bufs[r]) + slot_off : nullptr;
        flags.f[r] = r < n_ranks ? reinterpret_cast<uint32_t*>(peer_flags[r]) : nullptr;
    }
    finalize_top1_p2p_kernel<<<(n_pods + 31) / 32, 256, 0, s>>>(e->d_nodes, e->n_nodes, d_pods, n_pods, d_hint, hint_row_bytes(e), d_out, best, peers, n_ranks, my_rank, e->params,
                                                                flags, epoch, e->d_done + 1);
    CU(e, cudaGetLastError());
    merge_top1_wait_kernel<<<(n_pods + 127) / 128, 128, 0, s>>>(peers.p[my_rank], flags.f[my_rank], n_ranks, epoch, n_pods, d_merged);
    CU(e, cudaGetLastError());
    e->stats.kernel_launches += 2;
    return note_use(e, s);
}

// Latency path (per-pod Prioritize, and Schedule of a single workload): no copy operations at all.  Pod descriptors, hint rows
// and placements live in pinned host memory, which the kernel reads and writes directly over PCIe (unified addressing; up to
// four descriptors travel in the kernel parameters): ONE launch, the last CTA to finish turns the keys into placements and
// raises a flag in pinned memory that the host spins on.  n_pods <= 32.
static int latency_top1(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods, const uint8_t* hint, kgwe_placement* out_topk, cudaStream_t s) {
    const uint32_t row = hint_row_bytes(e);
    int rc;
    CU(e, e->h_pods.ensure(32)); CU(e, e->h_topk.ensure(32));
    if (hint) CU(e, e->h_hint.ensure((size_t)32 * row));
    memcpy(e->h_pods.p, pods, (size_t)n_pods * sizeof(kgwe_pod_req));
    if (hint) memcpy(e->h_hint.p, hint, (size_t)n_pods * row);
    rc = launch_score(e, e->h_pods.p, n_pods, hint ? e->h_hint.p : nullptr, nullptr, 1, e->h_topk.p, s, false, true);
    if (rc != KGWE_OK) return rc;
    // wait for the kernel's flag in pinned memory instead of a stream synchronisation; look at the
    // stream now and then so a failed launch cannot spin forever
    volatile unsigned int* flag = e->h_flag;
    const auto t0 = std::chrono::steady_clock::now();
    const char* trouble = nullptr; cudaError_t q = cudaSuccess;
    for (uint32_t spins = 0; *flag != e->flag_seq; ++spins) {
        if ((spins & 0x3fff) == 0x3fff) {
            q = cudaStreamQuery(s);
            if (q == cudaSuccess) { if (*flag != e->flag_seq) trouble = "kernel finished without raising its flag"; break; }
            if (q != cudaErrorNotReady) { trouble = cudaGetErrorString(q); break; }
            if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 5.0) { trouble = "no answer from the GPU after 5 s"; break; }
        }
    }
    if (trouble) {
        // leave the ticket counter of the fused tail at zero for the next call, whatever state the launch ended in
        cudaStreamSynchronize(s);
        cudaMemsetAsync(e->d_done, 0, sizeof(unsigned int), s);
        cudaStreamSynchronize(s);
        cudaGetLastError();
        return fail(e, KGWE_ERR_CUDA, "latency path: %s (flag %u, expected %u)", trouble, *flag, e->flag_seq);
    }
    rc = note_use(e, s);
    if (rc != KGWE_OK) return rc;
    memcpy(out_topk, e->h_topk.p, (size_t)n_pods * sizeof(kgwe_placement));
    e->stats.h2d_bytes += (uint64_t)n_pods * sizeof(kgwe_pod_req) + (hint ? (uint64_t)n_pods * row : 0);
    e->stats.d2h_bytes += (uint64_t)n_pods * sizeof(kgwe_placement);
    return KGWE_OK;
}

int kgwe_score_batch(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods, const uint8_t* hint,
                     kgwe_node_score* out_scores, uint32_t topk, kgwe_placement* out_topk) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (topk > KGWE_MAX_TOPK || (topk && !out_topk) || (n_pods && !pods)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    CallTimer timer(e, n_pods);
    int rc = validate_pods(e, pods, n_pods);
    if (rc != KGWE_OK) return rc;
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = e->stream;
    rc = flush_dirty(e, s);
    if (rc != KGWE_OK) return rc;
    if (n_pods == 0) return KGWE_OK;
    const uint32_t N = e->n_nodes;
    if (N == 0) {
        for (size_t i = 0; i < (size_t)n_pods * topk; ++i) { memset(&out_topk[i], 0, sizeof(kgwe_placement)); out_topk[i].node_idx = KGWE_NO_NODE; }
        return KGWE_OK;
    }
    const uint32_t row = hint_row_bytes(e);
    const bool need_matrix = out_scores || topk > 1;
    if (!need_matrix && topk == 1 && n_pods <= 32 && !(e->cfg.flags & KGWE_CFG_SUBSET_TABLES)) {
        rc = latency_top1(e, pods, n_pods, hint, out_topk, s);
        if (rc != KGWE_OK) return rc;
        timer.done(false);
        return KGWE_OK;
    }
    // pods per pass bounded by the score-matrix scratch
    uint32_t pass = n_pods;
    if (need_matrix) pass = (uint32_t)std::max<size_t>(1, std::min<size_t>(n_pods, score_scratch_bytes() / ((size_t)N * sizeof(kgwe_node_score))));
    CU(e, e->d_pods.ensure(pass)); CU(e, e->h_pods.ensure(pass));
    if (hint) { CU(e, e->d_hint.ensure((size_t)pass * row)); CU(e, e->h_hint.ensure((size_t)pass * row)); }
    if (topk) { CU(e, e->d_topk.ensure((size_t)pass * topk)); CU(e, e->h_topk.ensure((size_t)pass * topk)); }
    if (need_matrix) CU(e, e->d_scores.ensure((size_t)pass * N));
    double gpu_ms = 0.0;
    for (uint32_t p0 = 0; p0 < n_pods; p0 += pass) {
        const uint32_t np = std::min(pass, n_pods - p0);
        memcpy(e->h_pods.p, pods + p0, (size_t)np * sizeof(kgwe_pod_req));
        CU(e, cudaMemcpyAsync(e->d_pods.p, e->h_pods.p, (size_t)np * sizeof(kgwe_pod_req), cudaMemcpyHostToDevice, s));
        e->stats.h2d_bytes += (uint64_t)np * sizeof(kgwe_pod_req);
        if (hint) {
            memcpy(e->h_hint.p, hint + (size_t)p0 * row, (size_t)np * row);
            CU(e, cudaMemcpyAsync(e->d_hint.p, e->h_hint.p, (size_t)np * row, cudaMemcpyHostToDevice, s));
            e->stats.h2d_bytes += (uint64_t)np * row;
        }
        CU(e, cudaEventRecord(e->ev0, s));
        bool any_exh = false;
        for (uint32_t i = 0; i < np && !any_exh; ++i) any_exh = pods[p0 + i].flags & KGWE_POD_EXHAUSTIVE;
        rc = launch_score(e, e->d_pods.p, np, hint ? e->d_hint.p : nullptr, need_matrix ? e->d_scores.p : nullptr, topk, topk ? e->d_topk.p : nullptr, s, any_exh);
        if (rc != KGWE_OK) return rc;
        CU(e, cudaEventRecord(e->ev1, s));
        if (topk) {
            CU(e, cudaMemcpyAsync(e->h_topk.p, e->d_topk.p, (size_t)np * topk * sizeof(kgwe_placement), cudaMemcpyDeviceToHost, s));
            e->stats.d2h_bytes += (uint64_t)np * topk * sizeof(kgwe_placement);
        }
        if (out_scores) {
            CU(e, cudaMemcpyAsync(out_scores + (size_t)p0 * N, e->d_scores.p, (size_t)np * N * sizeof(kgwe_node_score), cudaMemcpyDeviceToHost, s));
            e->stats.d2h_bytes += (uint64_t)np * N * sizeof(kgwe_node_score);
        }
        rc = pull_counters(e, s);
        if (rc != KGWE_OK) return rc;
        CU(e, cudaStreamSynchronize(s));
        if (topk) memcpy(out_topk + (size_t)p0 * topk, e->h_topk.p, (size_t)np * topk * sizeof(kgwe_placement));
        fold_counters(e);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e->ev0, e->ev1) == cudaSuccess) gpu_ms += ms;
    }
    e->stats.last_call_gpu_ms = gpu_ms;
    timer.done(false);
    return KGWE_OK;
}

int kgwe_topk_merge_device(kgwe_engine* e, const kgwe_placement* d_lists, uint32_t n_lists, uint32_t n_pods, uint32_t topk,
                           kgwe_placement* d_out, void* cuda_stream) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (!d_lists || !d_out || topk == 0 || topk > KGWE_MAX_TOPK || n_lists == 0 || n_lists > 32) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    if (n_pods == 0) return KGWE_OK;
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : e->stream;
    topk_merge_kernel<<<(n_pods + 127) / 128, 128, 0, s>>>(d_lists, n_lists, n_pods, topk, d_out);
    CU(e, cudaGetLastError());
    e->stats.kernel_launches++;
    return KGWE_OK;
}

int kgwe_topk_merge_host(const kgwe_placement* lists, uint32_t n_lists, uint32_t n_pods, uint32_t topk, kgwe_placement* out) {
    if (!lists || !out || topk == 0 || topk > KGWE_MAX_TOPK || n_lists == 0 || n_lists > 32) return KGWE_ERR_INVALID_ARG;
    for (uint32_t p = 0; p < n_pods; ++p) {
        uint32_t head[32] = {0};
        for (uint32_t j = 0; j < topk; ++j) {
            int br = -1; int32_t bt = 0; uint32_t bn = 0;
            for (uint32_t r = 0; r < n_lists; ++r) {
                if (head[r] >= topk) continue;
                const kgwe_placement& c = lists[((size_t)r * n_pods + p) * topk + head[r]];
                if (c.node_idx == KGWE_NO_NODE) continue;
                if (br < 0 || c.total > bt || (c.total == bt && c.node_idx < bn)) { br = (int)r; bt = c.total; bn = c.node_idx; }
            }
            kgwe_placement& dst = out[(size_t)p * topk + j];
            if (br < 0) { memset(&dst, 0, sizeof dst); dst.node_idx = KGWE_NO_NODE; }
            else { dst = lists[((size_t)br * n_pods + p) * topk + head[br]]; head[br]++; }
        }
    }
    return KGWE_OK;
}

// Launch the stateful queue kernel (K3) for np pods and wait.  Decisions land in e->h_topk; the
// DEVICE image then carries every commit of the segment, the host mirror none yet.
// True when the image fits the tile form of the queue kernel (every node has its own lane group).
static bool queue_tile_form(const kgwe_engine* e) {
    const uint32_t n_tiles = (e->n_nodes + TILE_NODES - 1) / TILE_NODES;
    return n_tiles <= (uint32_t)(e->sm_count * e->sched_tile_blocks_per_sm);
}

// True when the image fits the cluster form (one lane per node inside one thread-block cluster, kgwe_queue.cuh).
static bool queue_cluster_form(const kgwe_engine* e) {
    return e->qc_max_ctas > 0 && e->n_nodes <= (uint32_t)e->qc_max_ctas * (uint32_t)QC_MAX_THREADS && e->max_total < (1ll << (32 - QC_NODE_BITS));
}

// gang_of (host, np entries, nullable): all-or-nothing gangs decided on the device (cluster and tile forms only).
static int run_queue_segment(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t np, const uint8_t* hint_rows, double* gpu_ms, const uint32_t* gang_of = nullptr,
                             const QcPeers* peers = nullptr) {
    cudaStream_t s = e->stream;
    const uint32_t N = e->n_nodes, row = hint_row_bytes(e);
    int rc = flush_dirty(e, s);
    if (rc != KGWE_OK) return rc;
    CU(e, e->d_pods.ensure(np)); CU(e, e->h_pods.ensure(np));
    CU(e, e->d_topk.ensure(np)); CU(e, e->h_topk.ensure(np));
    {
        const size_t had = e->d_best.cap;
        CU(e, e->d_best.ensure(np));
        if (e->d_best.cap != had) CU(e, cudaMemsetAsync(e->d_best.p, 0, e->d_best.cap * sizeof(unsigned long long), s));
        const size_t had2 = e->d_arrive.cap;
        CU(e, e->d_arrive.ensure(np));
        if (e->d_arrive.cap != had2) CU(e, cudaMemsetAsync(e->d_arrive.p, 0, e->d_arrive.cap * sizeof(unsigned int), s));
    }
    memcpy(e->h_pods.p, pods, (size_t)np * sizeof(kgwe_pod_req));
    CU(e, cudaMemcpyAsync(e->d_pods.p, e->h_pods.p, (size_t)np * sizeof(kgwe_pod_req), cudaMemcpyHostToDevice, s));
    e->stats.h2d_bytes += (uint64_t)np * sizeof(kgwe_pod_req);
    const uint8_t* d_hint = nullptr;
    if (hint_rows) {
        CU(e, e->d_hint.ensure((size_t)np * row)); CU(e, e->h_hint.ensure((size_t)np * row));
        memcpy(e->h_hint.p, hint_rows, (size_t)np * row);
        CU(e, cudaMemcpyAsync(e->d_hint.p, e->h_hint.p, (size_t)np * row, cudaMemcpyHostToDevice, s));
        e->stats.h2d_bytes += (uint64_t)np * row;
        d_hint = e->d_hint.p;
    }
    const uint32_t* d_gang = nullptr;
    if (gang_of) {
        CU(e, e->d_gang.ensure(np));
        CU(e, cudaMemcpyAsync(e->d_gang.p, gang_of, (size_t)np * sizeof(uint32_t), cudaMemcpyHostToDevice, s));      // (pageable source: the runtime stages it before returning)
        e->stats.h2d_bytes += (uint64_t)np * sizeof(uint32_t);
        d_gang = e->d_gang.p;
    }
    // persistent grid: as many CTAs as can be co-resident, no more than there are node tiles
    uint32_t grid = std::min<uint32_t>((uint32_t)(e->sm_count * e->sched_blocks_per_sm), (N + TILE_NODES - 1) / TILE_NODES);
    grid = std::max(grid, 1u);
    DevNode* d_nodes = e->d_nodes; uint32_t n_nodes = N; const kgwe_pod_req* d_pods = e->d_pods.p; uint32_t npods = np; uint32_t rowb = row;
    unsigned long long* best = e->d_best.p; unsigned int* arrive = e->d_arrive.p; kgwe_placement* d_out = e->d_topk.p; Params prm = e->params; unsigned long long* ctr = e->d_counters;
    void* args[] = {&d_nodes, &n_nodes, &d_pods, &npods, &d_hint, &rowb, &best, &arrive, &d_out, &prm, &ctr};
    void* targs[] = {&d_nodes, &n_nodes, &d_pods, &npods, &d_hint, &rowb, &best, &arrive, &d_out, &prm, &ctr, &d_gang};
    CU(e, cudaEventRecord(e->ev0, s));
    const uint32_t n_tiles = (N + TILE_NODES - 1) / TILE_NODES;
    if (queue_cluster_form(e)) {
        // the whole image inside one cluster: 128 lanes per CTA while 16 such CTAs hold it, else 640
        const uint32_t threads = N <= (uint32_t)e->qc_max_ctas * (uint32_t)QC_SMALL_THREADS ? (uint32_t)QC_SMALL_THREADS : (uint32_t)QC_MAX_THREADS;
        uint32_t ctas = 1;
        while (ctas < (uint32_t)e->qc_max_ctas && ctas * threads < N) ctas <<= 1;
        cudaLaunchConfig_t lc = {};
        lc.gridDim = dim3(ctas); lc.blockDim = dim3(threads); lc.dynamicSmemBytes = qc_layout(threads).total; lc.stream = s;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = ctas; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        lc.attrs = at; lc.numAttrs = 1;
        const QcPeers pq = peers ? *peers : QcPeers{};
        if (peers) {
            if (threads == (uint32_t)QC_SMALL_THREADS) CU(e, cudaLaunchKernelEx(&lc, schedule_cluster_kernel<QC_SMALL_THREADS, true>, d_nodes, n_nodes, d_pods, npods, d_hint, rowb, d_out, prm, ctr, d_gang, pq));
            else CU(e, cudaLaunchKernelEx(&lc, schedule_cluster_kernel<QC_MAX_THREADS, true>, d_nodes, n_nodes, d_pods, npods, d_hint, rowb, d_out, prm, ctr, d_gang, pq));
        } else if (threads == (uint32_t)QC_SMALL_THREADS) CU(e, cudaLaunchKernelEx(&lc, schedule_cluster_kernel<QC_SMALL_THREADS, false>, d_nodes, n_nodes, d_pods, npods, d_hint, rowb, d_out, prm, ctr, d_gang, pq));
        else CU(e, cudaLaunchKernelEx(&lc, schedule_cluster_kernel<QC_MAX_THREADS, false>, d_nodes, n_nodes, d_pods, npods, d_hint, rowb, d_out, prm, ctr, d_gang, pq));
    } else if (queue_tile_form(e)) {
        // every node gets its own lane group: tile form (shared-memory resident tables)
        CU(e, cudaLaunchCooperativeKernel((const void*)schedule_tile_kernel, dim3(n_tiles), dim3(SCORE_THREADS), targs, sched_tile_smem_bytes(), s));
    } else
    CU(e, cudaLaunchCooperativeKernel((const void*)schedule_kernel, dim3(grid), dim3(SCORE_THREADS), args, 0, s));
    if (queue_cluster_form(e)) {      // (the cluster form fills in EstimatedBandwidth itself and uses neither key table nor arrival counters)
        CU(e, cudaGetLastError());
        e->stats.kernel_launches += 1;
    } else {
        est_bandwidth_kernel<<<(np + 127) / 128, 128, 0, s>>>(e->d_nodes, N, e->cfg.node_base, e->d_topk.p, np);
        CU(e, cudaGetLastError());
        e->stats.kernel_launches += 2;
        CU(e, cudaMemsetAsync(e->d_best.p, 0, (size_t)np * sizeof(unsigned long long), s));
        CU(e, cudaMemsetAsync(e->d_arrive.p, 0, (size_t)np * sizeof(unsigned int), s));
    }
    CU(e, cudaEventRecord(e->ev1, s));
    CU(e, cudaMemcpyAsync(e->h_topk.p, e->d_topk.p, (size_t)np * sizeof(kgwe_placement), cudaMemcpyDeviceToHost, s));
    e->stats.d2h_bytes += (uint64_t)np * sizeof(kgwe_placement);
    rc = pull_counters(e, s);
    if (rc != KGWE_OK) return rc;
    CU(e, cudaStreamSynchronize(s));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e->ev0, e->ev1) == cudaSuccess) *gpu_ms += ms;
    fold_counters(e);
    if (e->h_counters[3] != e->watchdog_seen) {
        e->watchdog_seen = e->h_counters[3];
        return fail(e, KGWE_ERR_CUDA, "queue kernel: the exchange between the CTAs of the cluster timed out; decisions of this call are invalid");
    }
    return KGWE_OK;
}

// Mirror one device-side commit into the host ledger and the pinned image (no device traffic).
static void apply_commit(kgwe_engine* e, uint64_t wl_id, const kgwe_placement& d) {
    e->stats.scheduled++;
    if (d.topology >= 80) e->stats.topology_optimal++;
    const uint32_t local = d.node_idx - e->cfg.node_base;
    wait_copy(e);
    DevNode& rec = e->h_nodes[local];
    rec.allocated_mask |= d.gpu_mask; rec.n_allocs++;
    e->ledger[wl_id] = LedgerEntry{local, d.gpu_mask};
    mark_dirty_shadow(e, local);          // the queue kernel committed on the front image itself; the other image has to follow
}

static void no_node(kgwe_placement& p, uint8_t flags) { memset(&p, 0, sizeof p); p.node_idx = KGWE_NO_NODE; p.flags = flags; }

int kgwe_schedule_queue(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods, const uint8_t* hint, kgwe_placement* out) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if ((n_pods && !pods) || (n_pods && !out)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    CallTimer timer(e, n_pods);
    int rc = validate_pods(e, pods, n_pods);
    if (rc != KGWE_OK) return rc;
    CU(e, cudaSetDevice(e->cfg.device));
    if (n_pods == 0) return KGWE_OK;
    if (e->n_nodes == 0) {
        for (uint32_t i = 0; i < n_pods; ++i) { no_node(out[i], 0); e->stats.failed++; }
        return KGWE_OK;
    }
    if (n_pods == 1 && !(e->cfg.flags & KGWE_CFG_SUBSET_TABLES)) {
        // One Schedule() call (scheduler.go:114-179): score on the latency path (one launch, no copies), then record the
        // allocation on the host exactly as kgwe_alloc_commit would - the same UID overwrites its previous entry AFTER the
        // workload was scored against it (:699) - and let the next call publish the changed record(s).
        cudaStream_t s1 = e->stream;
        rc = flush_dirty(e, s1);
        if (rc != KGWE_OK) return rc;
        rc = latency_top1(e, pods, 1, hint, out, s1);
        if (rc != KGWE_OK) return rc;
        if (out[0].node_idx == KGWE_NO_NODE) e->stats.failed++;
        else {
            e->stats.scheduled++;
            if (out[0].topology >= 80) e->stats.topology_optimal++;
            wait_copy(e);
            ledger_apply(e, pods[0].wl_id, out[0].node_idx - e->cfg.node_base, out[0].gpu_mask, true);
        }
        e->stats.last_call_gpu_ms = 0.0;
        timer.done(true);
        return KGWE_OK;
    }
    const uint32_t row = hint_row_bytes(e);
    // A workload whose UID is already in the ledger (or repeats inside the queue) overwrites its
    // old entry AFTER it has been scored against it (scheduler.go:699).  Such a pod ends a
    // segment: the kernel runs up to and including it, then the host drops the old entry.
    double gpu_ms = 0.0;
    uint32_t p0 = 0;
    while (p0 < n_pods) {
        uint32_t p1 = p0; bool overwrite = false;
        {
            std::unordered_map<uint64_t, int> seen;
            while (p1 < n_pods) {
                const uint64_t id = pods[p1].wl_id;
                if (e->ledger.count(id) || seen.count(id)) { overwrite = true; ++p1; break; }
                seen[id] = 1; ++p1;
            }
        }
        const uint32_t np = p1 - p0;
        rc = run_queue_segment(e, pods + p0, np, hint ? hint + (size_t)p0 * row : nullptr, &gpu_ms);
        if (rc != KGWE_OK) return rc;
        for (uint32_t i = 0; i < np; ++i) {
            const kgwe_placement& d = e->h_topk.p[i];
            out[p0 + i] = d;
            if (d.node_idx == KGWE_NO_NODE) { e->stats.failed++; continue; }
            const bool last_overwrites = overwrite && i == np - 1;
            const auto prev = last_overwrites ? e->ledger.find(pods[p0 + i].wl_id) : e->ledger.end();
            // (a repeated id whose earlier occurrence found no node has nothing to overwrite)
            if (prev == e->ledger.end()) { apply_commit(e, pods[p0 + i].wl_id, d); continue; }
            // device already holds the new commit; bring the mirror level, then drop the old entry
            const uint32_t old_node = prev->second.node;
            e->ledger.erase(pods[p0 + i].wl_id);
            apply_commit(e, pods[p0 + i].wl_id, d);
            const uint32_t local = d.node_idx - e->cfg.node_base;
            for (uint32_t nd : {old_node, local}) { recount_node(e, nd); mark_dirty(e, nd); }
        }
        p0 = p1;
    }
    e->stats.last_call_gpu_ms = gpu_ms;
    timer.done(true);
    return KGWE_OK;
}

// The stateful queue over a node-sharded cluster, on the device: every rank (one engine = one GPU = one shard) makes this
// call with the same workloads; the queue kernels of the ranks exchange one 64-bit word per workload through peer-mapped
// memory (kgwe_queue.cuh) and the rank that owns the winning node commits it.
int kgwe_schedule_queue_p2p(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods, const uint8_t* hint,
                            const uint64_t* peer_recv_bufs, uint32_t n_ranks, uint32_t my_rank, uint32_t seq_base, kgwe_placement* out) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (!peer_recv_bufs || n_ranks < 2 || n_ranks > 8 || my_rank >= n_ranks || seq_base >= QC_SEQ_MOD || (n_pods && (!pods || !out)) || n_pods > (1u << 20))
        return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments (2..8 ranks, seq_base < 2^24 - 1, at most 2^20 workloads per call)");
    CallTimer timer(e, n_pods);
    int rc = validate_pods(e, pods, n_pods);
    if (rc != KGWE_OK) return rc;
    {
        std::unordered_map<uint64_t, int> seen;
        for (uint32_t i = 0; i < n_pods; ++i) {
            if (e->ledger.count(pods[i].wl_id) || seen.count(pods[i].wl_id)) return fail(e, KGWE_ERR_INVALID_ARG, "pod %u: workload id already scheduled (release it on every rank first)", i);
            seen[pods[i].wl_id] = 1;
        }
    }
    CU(e, cudaSetDevice(e->cfg.device));
    if (n_pods == 0) return KGWE_OK;
    if (e->n_nodes == 0) return fail(e, KGWE_ERR_INVALID_ARG, "the cluster image of this shard is empty");
    if (!queue_cluster_form(e) || e->max_total >= (1ll << 17) || (uint64_t)e->cfg.node_base + e->n_nodes > (1ull << 22))
        return fail(e, KGWE_ERR_TOO_LARGE, "the sharded queue needs a shard of at most %u nodes, TotalScore below 2^17 and global node indices below 2^22",
                    (unsigned)(e->qc_max_ctas * QC_MAX_THREADS));
    QcPeers pq{};
    for (uint32_t r = 0; r < n_ranks; ++r) pq.recv[r] = reinterpret_cast<unsigned long long*>(peer_recv_bufs[r]);
    pq.n_ranks = n_ranks; pq.my_rank = my_rank; pq.seq_base = seq_base;
    double gpu_ms = 0.0;
    rc = run_queue_segment(e, pods, n_pods, hint, &gpu_ms, nullptr, &pq);
    if (rc != KGWE_OK) return rc;
    for (uint32_t i = 0; i < n_pods; ++i) {
        const kgwe_placement& d = e->h_topk.p[i];
        out[i] = d;
        if (d.node_idx != KGWE_NO_NODE) apply_commit(e, pods[i].wl_id, d);      // (a workload placed on another rank, or nowhere, leaves this shard as it was)
    }
    e->stats.last_call_gpu_ms = gpu_ms;
    timer.done(true);
    return KGWE_OK;
}

// E5.  Speculative: run a window of the queue in one kernel; everything before the first gang
// with a failed member is final.  That gang is rejected, the device records touched from there
// on are restored from the host mirror (which never saw them), and the queue resumes after it.
int kgwe_schedule_gangs(kgwe_engine* e, const kgwe_pod_req* pods, uint32_t n_pods, const uint32_t* gang_of,
                        const uint8_t* hint, kgwe_placement* out) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (n_pods && (!pods || !out || !gang_of)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    CallTimer timer(e, n_pods);
    int rc = validate_pods(e, pods, n_pods);
    if (rc != KGWE_OK) return rc;
    {
        std::unordered_map<uint64_t, int> seen;
        for (uint32_t i = 0; i < n_pods; ++i) {
            if (i && gang_of[i] < gang_of[i - 1]) return fail(e, KGWE_ERR_INVALID_ARG, "gang_of must be non-decreasing (pod %u)", i);
            if (e->ledger.count(pods[i].wl_id) || seen.count(pods[i].wl_id)) return fail(e, KGWE_ERR_INVALID_ARG, "pod %u: workload id already scheduled", i);
            seen[pods[i].wl_id] = 1;
        }
    }
    CU(e, cudaSetDevice(e->cfg.device));
    if (n_pods == 0) return KGWE_OK;
    if (e->n_nodes == 0) {
        for (uint32_t i = 0; i < n_pods; ++i) { no_node(out[i], 0); e->stats.failed++; }
        return KGWE_OK;
    }
    const uint32_t row = hint_row_bytes(e);
    if ((queue_cluster_form(e) || queue_tile_form(e)) && !getenv("KGWE_GANG_WINDOW")) {      // (the variable also forces the window path below, for tests)
        // All-or-nothing on the device: the queue kernel itself takes a rejected gang's commits back (schedule_tile_kernel),
        // so the whole queue is one launch per segment; the host only turns the decisions of rejected gangs into
        // "no node" (+ KGWE_F_GANG_ROLLBACK for members that had found one) and mirrors the surviving commits.
        const uint32_t kSegment = 1u << 20;
        double gms = 0.0;
        uint32_t q0 = 0;
        while (q0 < n_pods) {
            uint32_t q1 = std::min(n_pods, q0 + kSegment);
            while (q1 < n_pods && gang_of[q1] == gang_of[q1 - 1]) ++q1;
            const uint32_t np = q1 - q0;
            rc = run_queue_segment(e, pods + q0, np, hint ? hint + (size_t)q0 * row : nullptr, &gms, gang_of + q0);
            if (rc != KGWE_OK) return rc;
            uint32_t g0 = 0;
            while (g0 < np) {
                uint32_t g1 = g0 + 1; bool all = e->h_topk.p[g0].node_idx != KGWE_NO_NODE;
                while (g1 < np && gang_of[q0 + g1] == gang_of[q0 + g0]) { all = all && e->h_topk.p[g1].node_idx != KGWE_NO_NODE; ++g1; }
                for (uint32_t i = g0; i < g1; ++i) {
                    const kgwe_placement& d = e->h_topk.p[i];
                    if (all) { out[q0 + i] = d; apply_commit(e, pods[q0 + i].wl_id, d); }
                    else { no_node(out[q0 + i], d.node_idx != KGWE_NO_NODE ? KGWE_F_GANG_ROLLBACK : 0); e->stats.failed++; }
                }
                g0 = g1;
            }
            q0 = q1;
        }
        e->stats.last_call_gpu_ms = gms;
        timer.done(true);
        return KGWE_OK;
    }
    // Images too large for the tile form of the queue kernel: speculative windows.
    // speculative window (pods per queue launch); KGWE_GANG_WINDOW shrinks it so tests can cross many window edges on small inputs
    static const uint32_t kWindow = [] { const char* v = getenv("KGWE_GANG_WINDOW"); const long x = v ? atol(v) : 0; return x > 0 ? (uint32_t)x : 4096u; }();
    double gpu_ms = 0.0;
    uint32_t p0 = 0;
    while (p0 < n_pods) {
        // window = whole gangs, at least one
        uint32_t p1 = std::min(n_pods, p0 + kWindow);
        while (p1 < n_pods && gang_of[p1] == gang_of[p1 - 1]) ++p1;
        const uint32_t np = p1 - p0;
        rc = run_queue_segment(e, pods + p0, np, hint ? hint + (size_t)p0 * row : nullptr, &gpu_ms);
        if (rc != KGWE_OK) return rc;
        uint32_t bad = np;                                   // first pod of the first failing gang
        for (uint32_t i = 0; i < np; ++i)
            if (e->h_topk.p[i].node_idx == KGWE_NO_NODE) { bad = i; while (bad > 0 && gang_of[p0 + bad - 1] == gang_of[p0 + i]) --bad; break; }
        for (uint32_t i = 0; i < bad; ++i) { out[p0 + i] = e->h_topk.p[i]; apply_commit(e, pods[p0 + i].wl_id, e->h_topk.p[i]); }
        if (bad == np) { p0 = p1; continue; }
        uint32_t gend = bad;
        while (gend < np && gang_of[p0 + gend] == gang_of[p0 + bad]) ++gend;
        for (uint32_t i = bad; i < np; ++i) {                // undo on the device whatever the kernel committed from here on
            const kgwe_placement& d = e->h_topk.p[i];
            if (d.node_idx != KGWE_NO_NODE) mark_dirty(e, d.node_idx - e->cfg.node_base);
            if (i < gend) { no_node(out[p0 + i], d.node_idx != KGWE_NO_NODE ? KGWE_F_GANG_ROLLBACK : 0); e->stats.failed++; }
        }
        p0 += gend;
    }
    e->stats.last_call_gpu_ms = gpu_ms;
    timer.done(true);
    return KGWE_OK;
}

int kgwe_placement_terms(kgwe_engine* e, const kgwe_placement* placements, uint32_t n, kgwe_side_terms* out) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (n && (!placements || !out)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    if (n == 0) return KGWE_OK;
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = e->stream;
    int rc = flush_dirty(e, s);
    if (rc != KGWE_OK) return rc;
    CU(e, e->d_topk.ensure(n)); CU(e, e->h_topk.ensure(n));
    static_assert(sizeof(kgwe_side_terms) <= sizeof(kgwe_placement), "terms are staged in the placement buffers");
    CU(e, e->d_scores.ensure(((size_t)n * sizeof(kgwe_side_terms) + sizeof(kgwe_node_score) - 1) / sizeof(kgwe_node_score)));
    memcpy(e->h_topk.p, placements, (size_t)n * sizeof(kgwe_placement));
    CU(e, cudaMemcpyAsync(e->d_topk.p, e->h_topk.p, (size_t)n * sizeof(kgwe_placement), cudaMemcpyHostToDevice, s));
    kgwe_side_terms* d_terms = reinterpret_cast<kgwe_side_terms*>(e->d_scores.p);
    side_terms_kernel<<<(n + 3) / 4, 128, 0, s>>>(e->d_nodes, e->n_nodes, e->cfg.node_base, e->d_topk.p, n, d_terms);
    CU(e, cudaGetLastError());
    e->stats.kernel_launches++;
    CU(e, cudaMemcpyAsync(e->h_topk.p, d_terms, (size_t)n * sizeof(kgwe_side_terms), cudaMemcpyDeviceToHost, s));
    CU(e, cudaStreamSynchronize(s));
    memcpy(out, e->h_topk.p, (size_t)n * sizeof(kgwe_side_terms));
    e->stats.h2d_bytes += (uint64_t)n * sizeof(kgwe_placement); e->stats.d2h_bytes += (uint64_t)n * sizeof(kgwe_side_terms);
    return KGWE_OK;
}

int kgwe_topology_hint_batch(kgwe_engine* e, const kgwe_hint_req* reqs, uint32_t n_reqs, kgwe_topology_hint* out) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (n_reqs && (!reqs || !out)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    if (n_reqs > 65535u) return fail(e, KGWE_ERR_TOO_LARGE, "at most 65535 hint requests per call");
    for (uint32_t i = 0; i < n_reqs; ++i)
        if (reqs[i].min_memory_gb < 0) return fail(e, KGWE_ERR_INVALID_ARG, "request %u: min_memory_gb < 0", i);
    if (n_reqs == 0) return KGWE_OK;
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = e->stream;
    int rc = flush_dirty(e, s);
    if (rc != KGWE_OK) return rc;
    static_assert(sizeof(kgwe_hint_req) == 16 && sizeof(kgwe_topology_hint) == 24, "hint ABI sizes");
    if (e->n_nodes == 0) {
        for (uint32_t i = 0; i < n_reqs; ++i) { memset(&out[i], 0, sizeof out[i]); out[i].node_idx = KGWE_NO_NODE; }
        return KGWE_OK;
    }
    const size_t bytes_in = (size_t)n_reqs * sizeof(kgwe_hint_req), off_out = (bytes_in + 255) & ~(size_t)255, bytes_out = (size_t)n_reqs * sizeof(kgwe_topology_hint);
    CU(e, e->d_hint.ensure(off_out + bytes_out)); CU(e, e->h_hint.ensure(off_out + bytes_out));
    {
        const size_t had = e->d_best.cap;
        CU(e, e->d_best.ensure(n_reqs));
        if (e->d_best.cap != had) CU(e, cudaMemsetAsync(e->d_best.p, 0, e->d_best.cap * sizeof(unsigned long long), s));
    }
    memcpy(e->h_hint.p, reqs, bytes_in);
    CU(e, cudaMemcpyAsync(e->d_hint.p, e->h_hint.p, bytes_in, cudaMemcpyHostToDevice, s));
    const kgwe_hint_req* d_reqs = reinterpret_cast<const kgwe_hint_req*>(e->d_hint.p);
    kgwe_topology_hint* d_out = reinterpret_cast<kgwe_topology_hint*>(e->d_hint.p + off_out);
    hint_score_kernel<<<dim3((e->n_nodes + 255) / 256, n_reqs), 256, 0, s>>>(e->d_nodes, e->n_nodes, e->cfg.node_base, d_reqs, e->d_best.p);
    CU(e, cudaGetLastError());
    hint_finalize_kernel<<<(n_reqs + 127) / 128, 128, 0, s>>>(e->d_nodes, e->n_nodes, e->cfg.node_base, d_reqs, n_reqs, e->d_best.p, d_out);
    CU(e, cudaGetLastError());
    e->stats.kernel_launches += 2;
    CU(e, cudaMemcpyAsync(e->h_hint.p + off_out, d_out, bytes_out, cudaMemcpyDeviceToHost, s));
    CU(e, cudaStreamSynchronize(s));
    memcpy(out, e->h_hint.p + off_out, bytes_out);
    e->stats.h2d_bytes += bytes_in; e->stats.d2h_bytes += bytes_out;
    return KGWE_OK;
}

int kgwe_topology_gauges(kgwe_engine* e, int32_t* out, uint32_t n_nodes) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (n_nodes && !out) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    if (n_nodes == 0) return KGWE_OK;
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = e->stream;
    int rc = flush_dirty(e, s);
    if (rc != KGWE_OK) return rc;
    const size_t bytes = (size_t)n_nodes * sizeof(int32_t);
    CU(e, e->d_hint.ensure(bytes)); CU(e, e->h_hint.ensure(bytes));
    gauge_kernel<<<(n_nodes + 255) / 256, 256, 0, s>>>(e->d_nodes, std::min(e->n_nodes, e->cfg.max_nodes), reinterpret_cast<int32_t*>(e->d_hint.p), n_nodes);
    CU(e, cudaGetLastError());
    e->stats.kernel_launches++;
    CU(e, cudaMemcpyAsync(e->h_hint.p, e->d_hint.p, bytes, cudaMemcpyDeviceToHost, s));
    CU(e, cudaStreamSynchronize(s));
    memcpy(out, e->h_hint.p, bytes);
    e->stats.d2h_bytes += bytes;
    return KGWE_OK;
}

// E2.  The MIG table is the caller's; the engine lends its device, stream and scratch.
int kgwe_mig_pack(kgwe_engine* e, kgwe_mig_gpu* gpus, uint32_t n_gpus, const uint8_t* profiles, uint32_t n_reqs, kgwe_mig_assignment* out) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if ((n_gpus && !gpus) || (n_reqs && (!profiles || !out))) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    if (n_reqs == 0 || n_gpus == 0) {
        for (uint32_t i = 0; i < n_reqs; ++i) { memset(&out[i], 0, sizeof out[i]); out[i].gpu_idx = KGWE_NO_NODE; out[i].profile = profiles[i]; }
        return KGWE_OK;
    }
    // tables up to MIG_TREE_MAX_GPUS: the tournament-tree kernel (one shared-memory read per request); larger ones: the
    // scan form (per-thread cached keys, block-wide min), up to what one CTA's shared memory holds
    const bool tree = n_gpus <= MIG_TREE_MAX_GPUS && mig_tree_smem(n_gpus) <= 220 * 1024;
    const size_t smem = tree ? mig_tree_smem(n_gpus)
                             : (size_t)MIG_GEOS * 256 + 36 * 4 + 10 * (size_t)MIG_THREADS * 4 + (size_t)MIG_THREADS * 2 + 2 * (size_t)n_gpus;
    if (smem > 220 * 1024 || n_gpus >= (1u << 24)) return fail(e, KGWE_ERR_TOO_LARGE, "kgwe_mig_pack: %u GPUs exceed the single-CTA table (max ~85k per call; split the table by GPU pool)", n_gpus);
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = e->stream;
    static_assert(sizeof(kgwe_mig_gpu) == 4 && sizeof(kgwe_mig_assignment) == 8 && sizeof(kgwe_mig_op) == 8, "MIG ABI sizes");
    const size_t bytes_g = (size_t)n_gpus * sizeof(kgwe_mig_gpu), bytes_o = (size_t)n_reqs * sizeof(kgwe_mig_assignment);
    // scratch carved from the byte buffers: [gpus][profiles][assignments]
    const size_t off_p = (bytes_g + 255) & ~(size_t)255, off_o = (off_p + n_reqs + 255) & ~(size_t)255, total = off_o + bytes_o;
    CU(e, e->d_hint.ensure(total)); CU(e, e->h_hint.ensure(total));
    memcpy(e->h_hint.p, gpus, bytes_g); memcpy(e->h_hint.p + off_p, profiles, n_reqs);
    CU(e, cudaMemcpyAsync(e->d_hint.p, e->h_hint.p, off_p + n_reqs, cudaMemcpyHostToDevice, s));
    CU(e, cudaFuncSetAttribute(tree ? (const void*)mig_pack_tree_kernel : (const void*)mig_pack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CU(e, cudaEventRecord(e->ev0, s));
    if (tree) mig_pack_tree_kernel<<<2, MIGT_THREADS, smem, s>>>(reinterpret_cast<kgwe_mig_gpu*>(e->d_hint.p), n_gpus, e->d_hint.p + off_p, n_reqs,
                                                                 reinterpret_cast<kgwe_mig_assignment*>(e->d_hint.p + off_o));
    else mig_pack_kernel<<<1, MIG_THREADS, smem, s>>>(reinterpret_cast<kgwe_mig_gpu*>(e->d_hint.p), n_gpus, e->d_hint.p + off_p, n_reqs,
                                                      reinterpret_cast<kgwe_mig_assignment*>(e->d_hint.p + off_o));
    CU(e, cudaGetLastError());
    CU(e, cudaEventRecord(e->ev1, s));
    e->stats.kernel_launches++;
    CU(e, cudaMemcpyAsync(e->h_hint.p, e->d_hint.p, bytes_g, cudaMemcpyDeviceToHost, s));
    CU(e, cudaMemcpyAsync(e->h_hint.p + off_o, e->d_hint.p + off_o, bytes_o, cudaMemcpyDeviceToHost, s));
    CU(e, cudaStreamSynchronize(s));
    memcpy(gpus, e->h_hint.p, bytes_g); memcpy(out, e->h_hint.p + off_o, bytes_o);
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e->ev0, e->ev1) == cudaSuccess) e->stats.last_call_gpu_ms = ms;
    e->stats.h2d_bytes += off_p + n_reqs; e->stats.d2h_bytes += bytes_g + bytes_o;
    return KGWE_OK;
}

int kgwe_mig_rebalance(kgwe_engine* e, const kgwe_mig_gpu* gpus, uint32_t n_gpus, const double* target,
                       kgwe_mig_op* ops, uint32_t ops_cap, uint32_t* n_ops, uint32_t* supply) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if ((n_gpus && !gpus) || !target || !n_ops || !supply || (ops_cap && !ops)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    double sum = 0.0;
    for (int p = 0; p < KGWE_MIG_PROFILES; ++p) { if (!(target[p] >= 0.0)) return fail(e, KGWE_ERR_INVALID_ARG, "target[%d] must be >= 0", p); sum += target[p]; }
    if (sum > 1.01) return fail(e, KGWE_ERR_INVALID_ARG, "profile distribution exceeds 100%%");      // mig_controller.go:263-270
    *n_ops = 0; for (int p = 0; p < KGWE_MIG_PROFILES; ++p) supply[p] = 0;
    if (n_gpus == 0) return KGWE_OK;
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = e->stream;
    const size_t bytes_g = (size_t)n_gpus * sizeof(kgwe_mig_gpu);
    const size_t off_r = (bytes_g + 255) & ~(size_t)255, off_c = off_r + (size_t)n_gpus * 4, off_o = (off_c + 8 + 255) & ~(size_t)255;
    const size_t total = off_o + (size_t)ops_cap * sizeof(kgwe_mig_op);
    CU(e, e->d_hint.ensure(total)); CU(e, e->h_hint.ensure(total));
    memcpy(e->h_hint.p, gpus, bytes_g);
    CU(e, cudaMemcpyAsync(e->d_hint.p, e->h_hint.p, bytes_g, cudaMemcpyHostToDevice, s));
    uint32_t* d_rank = reinterpret_cast<uint32_t*>(e->d_hint.p + off_r);
    uint32_t* d_cnt = reinterpret_cast<uint32_t*>(e->d_hint.p + off_c);
    mig_rank_kernel<<<1, 1024, 0, s>>>(reinterpret_cast<const kgwe_mig_gpu*>(e->d_hint.p), n_gpus, d_rank, d_cnt);
    CU(e, cudaGetLastError());
    uint32_t idle[2] = {0, 0};
    CU(e, cudaMemcpyAsync(e->h_hint.p + off_c, d_cnt, 8, cudaMemcpyDeviceToHost, s));
    CU(e, cudaStreamSynchronize(s));
    memcpy(idle, e->h_hint.p + off_c, 8);
    MigPlan plan;
    mig_make_plan(idle, target, plan, n_ops, supply);
    const uint32_t n_write = std::min(*n_ops, ops_cap);
    if (n_write) {
        mig_emit_kernel<<<(n_gpus + 255) / 256, 256, 0, s>>>(reinterpret_cast<const kgwe_mig_gpu*>(e->d_hint.p), n_gpus, d_rank, plan,
                                                            reinterpret_cast<kgwe_mig_op*>(e->d_hint.p + off_o), ops_cap);
        CU(e, cudaGetLastError());
        CU(e, cudaMemcpyAsync(e->h_hint.p + off_o, e->d_hint.p + off_o, (size_t)n_write * sizeof(kgwe_mig_op), cudaMemcpyDeviceToHost, s));
        CU(e, cudaStreamSynchronize(s));
        memcpy(ops, e->h_hint.p + off_o, (size_t)n_write * sizeof(kgwe_mig_op));
    }
    e->stats.kernel_launches += 2;
    return KGWE_OK;
}

int kgwe_mig_reconfigure(kgwe_engine* e, const kgwe_mig_gpu_state* gpus, uint32_t n_gpus, const double* target,
                         kgwe_mig_op* ops, uint32_t ops_cap, uint32_t* n_ops, uint32_t* supply) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if ((n_gpus && !gpus) || !target || !n_ops || !supply || (ops_cap && !ops)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    double sum = 0.0;
    for (int p = 0; p < KGWE_MIG_PROFILES; ++p) { if (!(target[p] >= 0.0)) return fail(e, KGWE_ERR_INVALID_ARG, "target[%d] must be >= 0", p); sum += target[p]; }
    if (sum > 1.01) return fail(e, KGWE_ERR_INVALID_ARG, "profile distribution exceeds 100%%");      // mig_controller.go:263-270
    for (uint32_t g = 0; g < n_gpus; ++g)
        if ((gpus[g].in_use & ~gpus[g].created) || (gpus[g].created >> 14)) return fail(e, KGWE_ERR_INVALID_ARG, "GPU %u: in_use must be a subset of created, both within the 14 legal placements", g);
    *n_ops = 0; for (int p = 0; p < KGWE_MIG_PROFILES; ++p) supply[p] = 0;
    if (n_gpus == 0) return KGWE_OK;
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = e->stream;
    static_assert(sizeof(kgwe_mig_gpu_state) == 8, "MIG ABI sizes");
    const size_t bytes_g = (size_t)n_gpus * sizeof(kgwe_mig_gpu_state);
    const size_t off_r = (bytes_g + 255) & ~(size_t)255, off_c = off_r + (size_t)n_gpus * 4, off_o = (off_c + 64 + 255) & ~(size_t)255;
    const size_t total = off_o + (size_t)ops_cap * sizeof(kgwe_mig_op);
    CU(e, e->d_hint.ensure(total)); CU(e, e->h_hint.ensure(total));
    memcpy(e->h_hint.p, gpus, bytes_g);
    CU(e, cudaMemcpyAsync(e->d_hint.p, e->h_hint.p, bytes_g, cudaMemcpyHostToDevice, s));
    const kgwe_mig_gpu_state* d_g = reinterpret_cast<const kgwe_mig_gpu_state*>(e->d_hint.p);
    uint32_t* d_rank = reinterpret_cast<uint32_t*>(e->d_hint.p + off_r);
    uint32_t* d_cnt = reinterpret_cast<uint32_t*>(e->d_hint.p + off_c);        // [0..1] reconfigurable per class, [2] ops, [4..13] supply
    mig_state_rank_kernel<<<1, 1024, 0, s>>>(d_g, n_gpus, d_rank, d_cnt);
    CU(e, cudaGetLastError());
    CU(e, cudaMemcpyAsync(e->h_hint.p + off_c, d_cnt, 8, cudaMemcpyDeviceToHost, s));
    CU(e, cudaStreamSynchronize(s));
    uint32_t counts[2];
    memcpy(counts, e->h_hint.p + off_c, 8);
    MigRanks plan;
    mig_split(counts, target, plan);
    mig_reconf_kernel<<<1, 1024, 0, s>>>(d_g, n_gpus, d_rank, plan, reinterpret_cast<kgwe_mig_op*>(e->d_hint.p + off_o), ops_cap, d_cnt + 2, d_cnt + 4);
    CU(e, cudaGetLastError());
    CU(e, cudaMemcpyAsync(e->h_hint.p + off_c, d_cnt, 64, cudaMemcpyDeviceToHost, s));
    CU(e, cudaStreamSynchronize(s));
    uint32_t res[16];
    memcpy(res, e->h_hint.p + off_c, 64);
    *n_ops = res[2];
    for (int p = 0; p < KGWE_MIG_PROFILES; ++p) supply[p] = res[4 + p];
    const uint32_t n_write = std::min(*n_ops, ops_cap);
    if (n_write) {
        CU(e, cudaMemcpyAsync(e->h_hint.p + off_o, e->d_hint.p + off_o, (size_t)n_write * sizeof(kgwe_mig_op), cudaMemcpyDeviceToHost, s));
        CU(e, cudaStreamSynchronize(s));
        memcpy(ops, e->h_hint.p + off_o, (size_t)n_write * sizeof(kgwe_mig_op));
    }
    e->stats.kernel_launches += 2;
    return KGWE_OK;
}

namespace {
struct SnapHeader { uint32_t magic, version, n_nodes, n_present; uint64_t n_ledger; uint32_t record_bytes, entry_bytes; };
struct SnapEntry { uint64_t wl_id; uint32_t node; uint32_t mask; };
constexpr uint32_t kSnapMagic = 0x4B475745u, kSnapVersion = 1;
uint64_t snapshot_bytes(const kgwe_engine* e) { return sizeof(SnapHeader) + (uint64_t)e->n_nodes * sizeof(DevNode) + (uint64_t)e->ledger.size() * sizeof(SnapEntry); }
}  // namespace

int kgwe_snapshot_size(kgwe_engine* e, uint64_t* bytes) {
    if (!e || !bytes) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    *bytes = snapshot_bytes(e);
    return KGWE_OK;
}

int kgwe_snapshot_save(kgwe_engine* e, void* buf, uint64_t capacity, uint64_t* written) {
    if (!e || !buf || !written) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    const uint64_t need = snapshot_bytes(e);
    if (capacity < need) return fail(e, KGWE_ERR_TOO_LARGE, "snapshot needs %llu bytes, buffer has %llu", (unsigned long long)need, (unsigned long long)capacity);
    // the pinned mirror is authoritative: every commit of the queue kernels is mirrored before a call returns
    unsigned char* p = static_cast<unsigned char*>(buf);
    SnapHeader h{kSnapMagic, kSnapVersion, e->n_nodes, e->n_present, (uint64_t)e->ledger.size(), (uint32_t)sizeof(DevNode), (uint32_t)sizeof(SnapEntry)};
    memcpy(p, &h, sizeof h); p += sizeof h;
    memcpy(p, e->h_nodes, (size_t)e->n_nodes * sizeof(DevNode)); p += (size_t)e->n_nodes * sizeof(DevNode);
    std::vector<SnapEntry> entries; entries.reserve(e->ledger.size());
    for (auto& kv : e->ledger) entries.push_back(SnapEntry{kv.first, kv.second.node, kv.second.mask});
    std::sort(entries.begin(), entries.end(), [](const SnapEntry& a, const SnapEntry& b) { return a.wl_id < b.wl_id; });   // deterministic blob
    memcpy(p, entries.data(), entries.size() * sizeof(SnapEntry));
    *written = need;
    return KGWE_OK;
}

int kgwe_snapshot_load(kgwe_engine* e, const void* buf, uint64_t len) {
    if (!e || !buf) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    SnapHeader h;
    if (len < sizeof h) return fail(e, KGWE_ERR_INVALID_ARG, "snapshot too short");
    memcpy(&h, buf, sizeof h);
    if (h.magic != kSnapMagic || h.version != kSnapVersion || h.record_bytes != sizeof(DevNode) || h.entry_bytes != sizeof(SnapEntry))
        return fail(e, KGWE_ERR_INVALID_ARG, "not a kgwe snapshot of this version");
    if (h.n_nodes > e->cfg.max_nodes) return fail(e, KGWE_ERR_OUT_OF_RANGE, "snapshot has %u nodes, engine holds %u", h.n_nodes, e->cfg.max_nodes);
    if (len != sizeof h + (uint64_t)h.n_nodes * sizeof(DevNode) + h.n_ledger * sizeof(SnapEntry)) return fail(e, KGWE_ERR_INVALID_ARG, "snapshot length mismatch");
    const unsigned char* p = static_cast<const unsigned char*>(buf) + sizeof h;
    const unsigned char* pl = p + (size_t)h.n_nodes * sizeof(DevNode);
    // check the whole blob before touching the engine: a rejected snapshot leaves the state as it was
    for (uint32_t i = 0; i < h.n_nodes; ++i) {
        DevNode rec; memcpy(&rec, p + (size_t)i * sizeof(DevNode), sizeof rec);
        if (rec.n_gpus > KGWE_MAX_GPUS) return fail(e, KGWE_ERR_INVALID_ARG, "snapshot record %u: n_gpus %u", i, rec.n_gpus);
        for (int g = 0; g < KGWE_MAX_GPUS; ++g) {
            if (!(rec.pcie_bw[g] >= 0.0) || !(rec.pcie_bw[g] <= KGWE_MAX_BW_GBPS) || !std::isfinite(rec.util[g])) return fail(e, KGWE_ERR_INVALID_ARG, "snapshot record %u: GPU %d out of range", i, g);
            for (int c = 0; c < KGWE_MAX_GPUS; ++c)
                if (!(rec.bw_in[g][c] >= 0.0) || !(rec.bw_in[g][c] <= KGWE_MAX_BW_GBPS)) return fail(e, KGWE_ERR_INVALID_ARG, "snapshot record %u: link bandwidth out of range", i);
        }
    }
    for (uint64_t i = 0; i < h.n_ledger; ++i) {
        SnapEntry en; memcpy(&en, pl + i * sizeof(SnapEntry), sizeof en);
        if (en.node >= e->cfg.max_nodes) return fail(e, KGWE_ERR_INVALID_ARG, "snapshot ledger entry names node %u", en.node);
    }
    const uint32_t old_n = e->n_nodes;
    wait_copy(e);
    memcpy(e->h_nodes, p, (size_t)h.n_nodes * sizeof(DevNode));
    for (uint32_t i = h.n_nodes; i < old_n; ++i) memset(&e->h_nodes[i], 0, sizeof(DevNode));
    e->ledger.clear();
    for (uint64_t i = 0; i < h.n_ledger; ++i) {
        SnapEntry en; memcpy(&en, pl + i * sizeof(SnapEntry), sizeof en);
        e->ledger[en.wl_id] = LedgerEntry{en.node, (uint8_t)en.mask};
    }
    // everything that can be derived is derived again, never trusted from a blob: masks cut to n_gpus, the whole-number and
    // uniform-link flags from the bandwidths, the ledger projection from the blob's own ledger
    for (uint32_t i = 0; i < h.n_nodes; ++i) {
        DevNode& r = e->h_nodes[i];
        if (!(r.flags & NODE_PRESENT)) { memset(&r, 0, sizeof r); continue; }
        const uint8_t gm = (uint8_t)((1u << r.n_gpus) - 1u);
        for (int g = 0; g < KGWE_MAX_GPUS; ++g) {
            r.act_in[g] &= gm; r.present[g] &= gm;
            if (g >= r.n_gpus) { r.act_in[g] = 0; r.present[g] = 0; }
            r.act_in[g] &= (uint8_t)~(1u << g); r.present[g] &= (uint8_t)~(1u << g);
        }
        r.healthy_mask &= gm; r.mig_enabled_mask &= gm; r.mig_has_free_mask &= r.mig_enabled_mask;
        r.flags = NODE_PRESENT | (record_is_integral(r) ? NODE_INTEGRAL : 0);
        set_uniform_unit(r);
        r.allocated_mask = 0; r.n_allocs = 0;
    }
    for (auto& kv : e->ledger) {
        DevNode& r = e->h_nodes[kv.second.node];
        if (r.flags & NODE_PRESENT) { r.allocated_mask |= kv.second.mask; r.n_allocs++; }       // (entries naming an absent record wait for the node to come back, see upsert_one)
    }
    e->n_nodes = std::max(h.n_nodes, old_n);          // high-water mark never shrinks (absent records score as absent)
    e->n_present = 0; e->n_nonuniform = 0; e->n_fractional = 0;
    for (uint32_t i = 0; i < e->n_nodes; ++i) { if (e->h_nodes[i].flags & NODE_PRESENT) e->n_present++; e->n_nonuniform += rec_nonuniform(e->h_nodes[i]); e->n_fractional += rec_fractional(e->h_nodes[i]); mark_dirty(e, i); }
    return KGWE_OK;
}

int kgwe_node_read(kgwe_engine* e, uint32_t idx, kgwe_node_image* out) {
    if (!e || !out) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if (idx >= e->cfg.max_nodes) return fail(e, KGWE_ERR_OUT_OF_RANGE, "node %u out of range", idx);
    CU(e, cudaSetDevice(e->cfg.device));
    int rc = flush_dirty(e, e->stream);
    if (rc != KGWE_OK) return rc;
    DevNode rec;
    CU(e, cudaMemcpyAsync(&rec, e->d_nodes + idx, sizeof rec, cudaMemcpyDeviceToHost, e->stream));
    CU(e, cudaStreamSynchronize(e->stream));
    to_abi_record(rec, *out);
    return KGWE_OK;
}

int kgwe_get_stats(kgwe_engine* e, kgwe_stats* out) {
    if (!e || !out) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    // K1 timings of launches that have finished (no waiting: unfinished ones stay for the next poll)
    int kept = 0;
    for (int i = 0; i < e->timers_pending; ++i) {
        float ms = 0.f;
        if (kept == 0 && cudaEventQuery(e->t1[i]) == cudaSuccess) {
            if (cudaEventElapsedTime(&ms, e->t0[i], e->t1[i]) == cudaSuccess) { e->stats.score_kernel_ms += ms; e->stats.score_kernel_timed++; }
        } else {
            std::swap(e->t0[kept], e->t0[i]); std::swap(e->t1[kept], e->t1[i]); ++kept;
        }
    }
    e->timers_pending = kept;
    // Device counters are monotonic, so they can be read on a stream of their own while kernels run: no device-wide
    // synchronisation; work still in flight is counted by a later poll.
    if (cudaSetDevice(e->cfg.device) == cudaSuccess &&
        cudaMemcpyAsync(e->h_counters, e->d_counters, kCounters * sizeof(unsigned long long), cudaMemcpyDeviceToHost, e->qstream) == cudaSuccess &&
        cudaStreamSynchronize(e->qstream) == cudaSuccess)
        fold_counters(e);
    else cudaGetLastError();
    *out = e->stats;
    out->n_nodes = e->n_nodes;
    out->n_allocations = (uint32_t)e->ledger.size();
    return KGWE_OK;
}

const char* kgwe_last_error(kgwe_engine* e) { (void)e; return tl_last_error.c_str(); }

}  // extern "C"

// =======================================================================================
// kgwe_score_top1_noalloc — see include/kgwe_engine.h.  Same kernels as the top-1 path of
// kgwe_score_batch_device, but the arg-max key table is this fixed array instead of the
// engine's growable cudaMalloc'd buffer, so nothing in the call can allocate.
//
// g_noalloc_keys is a device array: in host code its name is only a handle for the
// runtime, not an address a kernel can use.  The usable device address is looked up once
// per device (with that device current) via cudaGetSymbolAddress and cached.
static __device__ __align__(256) unsigned long long g_noalloc_keys[KGWE_NOALLOC_MAX_PODS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned long long* g_noalloc_ptr[kMaxDevices] = {};

// Called with the engine's device current.  First call per device resolves the address.
int noalloc_table(kgwe_engine* e, unsigned long long** out) {
    const int dev = e->cfg.device;
    if (dev < 0 || dev >= kMaxDevices) return fail(e, KGWE_ERR_INVALID_ARG, "device ordinal %d not supported by the no-alloc path", dev);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        cudaError_t rc = cudaGetSymbolAddress(&p, g_noalloc_keys);
        if (rc != cudaSuccess || !p) return fail(e, KGWE_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_keys) on device %d: %s", dev, cudaGetErrorString(rc));
        g_noalloc_ptr[dev] = static_cast<unsigned long long*>(p);
    }
    *out = g_noalloc_ptr[dev];
    return KGWE_OK;
}
}  // namespace

extern "C" int kgwe_score_top1_noalloc(kgwe_engine* e, const kgwe_pod_req* d_pods, uint32_t n_pods, const uint8_t* d_hint,
                                       kgwe_node_score* d_out, kgwe_placement* d_top1, void* cuda_stream) {
    if (!e) return KGWE_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    if ((n_pods && !d_pods) || (n_pods && !d_top1)) return fail(e, KGWE_ERR_INVALID_ARG, "bad arguments");
    if (n_pods > KGWE_NOALLOC_MAX_PODS)
        return fail(e, KGWE_ERR_TOO_LARGE, "kgwe_score_top1_noalloc: %u pods exceed the fixed key table of %u (KGWE_NOALLOC_MAX_PODS); split the batch", n_pods, KGWE_NOALLOC_MAX_PODS);
    CU(e, cudaSetDevice(e->cfg.device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : e->stream;
    unsigned long long* keys = nullptr;
    int rc = noalloc_table(e, &keys);
    if (rc != KGWE_OK) return rc;
    rc = flush_dirty(e, s);                 // H2D copies of pending upserts only; allocates nothing
    if (rc != KGWE_OK) return rc;
    if (n_pods == 0) return KGWE_OK;
    if (e->n_nodes == 0) return fail(e, KGWE_ERR_INVALID_ARG, "kgwe_score_top1_noalloc: the cluster image is empty");
    // the table still holds whatever an earlier (possibly failed) call left: clear our slots
    CU(e, cudaMemsetAsync(keys, 0, (size_t)n_pods * sizeof(unsigned long long), s));
    rc = launch_k1(e, d_pods, n_pods, d_hint, d_out, keys, s, false, FusedTail{}, false);
    if (rc != KGWE_OK) return rc;
    finalize_top1_kernel<<<(n_pods + 31) / 32, 256, 0, s>>>(e->d_nodes, e->n_nodes, d_pods, n_pods, d_hint, hint_row_bytes(e), d_out, keys, d_top1, 1, e->params);
    CU(e, cudaGetLastError());
    e->stats.kernel_launches++;
    return note_use(e, s);
}
