// kgwe_mig.cuh — EXTENSION E2: MIG slice bin-packing and the rebalancing pass (see the spec in
// include/kgwe_engine.h; the reference only has stubs, src/sharing/mig_controller.go:340-504).
#pragma once
#include "kgwe_device.cuh"

namespace kgwe {

constexpr int MIG_GEOS = 5;                       // 1g 2g 3g 4g 7g
__host__ __device__ inline int mig_geo(int profile) { return (profile - 1) >> 1; }
__host__ __device__ inline int mig_cls(int profile) { return (profile - 1) & 1; }
__host__ __device__ inline int mig_size(int geo) { return geo == 0 ? 1 : geo == 1 ? 2 : geo == 4 ? 8 : 4; }
__host__ __device__ inline int mig_nstarts(int geo) { return geo == 0 ? 7 : geo == 1 ? 3 : geo == 2 ? 2 : 1; }
__host__ __device__ inline int mig_start(int geo, int i) { return geo == 0 ? i : geo == 1 ? 2 * i : geo == 2 ? 4 * i : 0; }
__host__ __device__ inline uint32_t mig_mask(int geo, int i) { return ((1u << mig_size(geo)) - 1u) << mig_start(geo, i); }

// Best placement of geometry `geo` on a GPU with occupancy `occ`: (cost << 3) | start, 0xFF if none.
// cost = number of still-free legal placements (all 14) that the placement overlaps.
__host__ __device__ inline uint32_t mig_best(uint32_t occ, int geo) {
    uint32_t best = 0xFF;
    for (int i = 0; i < mig_nstarts(geo); ++i) {
        const uint32_t m = mig_mask(geo, i);
        if (occ & m) continue;
        uint32_t cost = 0;
        for (int g2 = 0; g2 < MIG_GEOS; ++g2)
            for (int i2 = 0; i2 < mig_nstarts(g2); ++i2) {
                const uint32_t q = mig_mask(g2, i2);
                if (!(occ & q) && (q & m)) ++cost;
            }
        const uint32_t key = (cost << 3) | (uint32_t)mig_start(geo, i);
        if (key < best) best = key;
    }
    return best;
}

// K4: one persistent CTA.  Per-GPU state words (enabled/class << 8 | occupancy) and the
// (geometry, occupancy) -> best-placement table live in shared memory.  Thread t owns GPUs t, t+1024, ...
// and caches, per (class, geometry), the best key among them; a commit only invalidates the owner thread's
// five cached keys of that class, so a request costs one cached read per thread (plus a 10-GPU rescan by
// the few threads whose cache is stale), a block-wide min of cost << 27 | gpu << 3 | start, and the commit.
constexpr int MIG_THREADS = 1024;
__global__ void __launch_bounds__(MIG_THREADS, 1)
mig_pack_kernel(kgwe_mig_gpu* __restrict__ gpus, uint32_t n_gpus, const uint8_t* __restrict__ profiles, uint32_t n_reqs,
                kgwe_mig_assignment* __restrict__ out) {
    extern __shared__ unsigned char msm[];
    uint8_t* lut = msm;                                                             // [5][256]
    uint32_t* wmin = reinterpret_cast<uint32_t*>(msm + MIG_GEOS * 256);             // [32] + [1] winner
    uint32_t* cache = wmin + 36;                                                    // [10][1024] best key per (class*5 + geometry, thread)
    uint16_t* stale = reinterpret_cast<uint16_t*>(cache + 10 * MIG_THREADS);        // [1024] bit ci: cache[ci][t] must be recomputed
    uint16_t* state = stale + MIG_THREADS;                                          // [n_gpus]
    __shared__ uint8_t req_stage[MIG_THREADS];        // the next 1024 request profiles
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < MIG_GEOS * 256; i += MIG_THREADS) lut[i] = (uint8_t)mig_best((uint32_t)(i & 255), i >> 8);
    for (uint32_t g = tid; g < n_gpus; g += MIG_THREADS) {
        const kgwe_mig_gpu v = gpus[g];
        state[g] = (uint16_t)(((v.enabled ? 1 : 0) | ((v.cls & 1) << 1)) << 8 | v.occ);
    }
    stale[tid] = 0x3FF;
    __syncthreads();
    for (uint32_t r = 0; r < n_reqs; ++r) {
        if ((r & (MIG_THREADS - 1)) == 0) {           // refill the staged profiles (the loop's trailing barrier orders this)
            if (r + tid < n_reqs) req_stage[tid] = profiles[r + tid];
            __syncthreads();
        }
        const int profile = req_stage[r & (MIG_THREADS - 1)];
        uint32_t key = 0xFFFFFFFFu;
        if (profile >= 1 && profile <= KGWE_MIG_PROFILES) {
            const int geo = mig_geo(profile), ci = mig_cls(profile) * MIG_GEOS + geo;
            if (stale[tid] >> ci & 1) {
                const uint32_t want = 1u | ((uint32_t)mig_cls(profile) << 1);
                const uint8_t* row = lut + geo * 256;
                uint32_t k1 = 0xFFFFFFFFu;
                for (uint32_t g = tid; g < n_gpus; g += MIG_THREADS) {
                    const uint32_t st = state[g];
                    const uint32_t b = row[st & 0xff];
                    const uint32_t k2 = ((b >> 3) << 27) | (g << 3) | (b & 7);
                    if ((st >> 8) == want && b != 0xFF && k2 < k1) k1 = k2;
                }
                cache[ci * MIG_THREADS + tid] = k1;
                stale[tid] = (uint16_t)(stale[tid] & ~(1u << ci));
            }
            key = cache[ci * MIG_THREADS + tid];
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) { const uint32_t o = __shfl_xor_sync(0xffffffffu, key, off); key = o < key ? o : key; }
        if (lane == 0) wmin[warp] = key;
        __syncthreads();
        if (warp == 0) {
            uint32_t k2 = wmin[lane];
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) { const uint32_t o = __shfl_xor_sync(0xffffffffu, k2, off); k2 = o < k2 ? o : k2; }
            if (lane == 0) {
                kgwe_mig_assignment a;
                if (k2 == 0xFFFFFFFFu) { a.gpu_idx = KGWE_NO_NODE; a.start = a.size = a.cost = 0; a.profile = (uint8_t)profile; }
                else {
                    const uint32_t g = (k2 >> 3) & 0xFFFFFFu; const int geo = mig_geo(profile);
                    a.gpu_idx = g; a.start = (uint8_t)(k2 & 7); a.size = (uint8_t)mig_size(geo); a.cost = (uint8_t)(k2 >> 27); a.profile = (uint8_t)profile;
                    state[g] = (uint16_t)(state[g] | (((1u << a.size) - 1u) << a.start));
                    const uint32_t owner = g & (MIG_THREADS - 1);
                    stale[owner] = (uint16_t)(stale[owner] | (0x1Fu << (mig_cls(profile) * MIG_GEOS)));    // that class's five geometries
                }
                out[r] = a;
            }
        }
        __syncthreads();
    }
    for (uint32_t g = tid; g < n_gpus; g += MIG_THREADS) gpus[g].occ = (uint8_t)(state[g] & 0xff);
}

// K4, tournament-tree form (tables of up to MIG_TREE_MAX_GPUS GPUs): requests of one GPU class are inherently sequential - each one
// must see the commits before it - and ONE CTA per class serves them, so what counts is the length of the per-request instruction chain (a warp
// issues in order), not the width of the machine.  Per (class, geometry) a 32-ary min tree over the GPUs keeps the best
// placement key of every group of 32 GPUs, of every group of 32 groups, and at the root: a request is ONE shared-memory read
// of its root, and its commit only repairs the path above the one GPU that changed - three warp-wide REDUX.MIN per tree, one
// value per lane and level, the five trees of the GPU's class repaired by five warps side by side.  The three rows a warp
// needs are loaded up front (their addresses only depend on the GPU); each level's fresh minimum replaces the stale entry of
// the next row in a register.  The key carries the winner's occupancy (cost << 27 | gpu << 11 | occ << 3 | start), so the
// commit needs no look-up of the GPU's state.  Leaves: one byte per (geometry, GPU), class << 7 | cost << 3 | start, 0x7F = no fit.
// (A two-level tree with several values per lane was measured slower - 20 vs 12 ms for 50,000 requests on 10,000 GPUs: the
// extra per-lane instructions cost more than the REDUX they save.)
constexpr int MIGT_THREADS = 160;                  // five warps: warp w repairs the tree of geometry w
constexpr uint32_t MIG_TREE_MAX_GPUS = 24576;      // 3 levels of 32 cover 32,768; shared memory (8.3 B per GPU) sets this lower
constexpr uint32_t MIG_KEY_NONE = 0xFFFFFFFFu;
__host__ __device__ inline size_t mig_tree_smem(uint32_t n_gpus) {
    const size_t n_pad = (n_gpus + 31) & ~31u, n1 = n_pad / 32, n2 = (n1 + 31) / 32;
    return MIG_GEOS * 256 + 2 * n_pad + MIG_GEOS * n_pad + 10 * 4 * (n1 + n2 * 32 + 1) + 1024 + 128;
}
__device__ __forceinline__ uint32_t mig_leaf_key(uint32_t leaf, uint32_t cls, uint32_t g, uint32_t occ) {
    return ((leaf >> 7) == cls && (leaf & 0x7Fu) != 0x7Fu) ? ((((leaf >> 3) & 0xFu) << 27) | (g << 11) | ((occ & 0xffu) << 3) | (leaf & 7u)) : MIG_KEY_NONE;
}
__global__ void __launch_bounds__(MIGT_THREADS, 1)
mig_pack_tree_kernel(kgwe_mig_gpu* __restrict__ gpus, uint32_t n_gpus, const uint8_t* __restrict__ profiles, uint32_t n_reqs,
                     kgwe_mig_assignment* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char msm[];
    const uint32_t n_pad = (n_gpus + 31u) & ~31u, n1 = n_pad / 32u, n2 = (n1 + 31u) / 32u;
    uint32_t* L1 = reinterpret_cast<uint32_t*>(msm);                        // [10][n1]
    uint32_t* L2 = L1 + 10 * n1;                                            // [10][32] (padded with MIG_KEY_NONE; n2 <= 32)
    uint32_t* root = L2 + 10 * 32;                                          // [10] (+ pad)
    uint16_t* state = reinterpret_cast<uint16_t*>(root + 16);               // [n_pad]: (enabled | cls << 1) << 8 | occ
    uint8_t* leaf = reinterpret_cast<uint8_t*>(state + n_pad);              // [5][n_pad]
    uint8_t* lut = leaf + MIG_GEOS * n_pad;                                 // [5][256]
    uint8_t* req_stage = lut + MIG_GEOS * 256;                              // [1024]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // Two CTAs, one per GPU class: a request only ever touches GPUs (and trees) of its profile's class, so the two classes are
    // independent request streams.  Each CTA walks the whole stream, serves the requests of its class in order and steps over the
    // others without a barrier (CTA 0 also answers requests with an unknown profile).
    const uint32_t my_cls = blockIdx.x;
    for (int i = tid; i < MIG_GEOS * 256; i += MIGT_THREADS) lut[i] = (uint8_t)mig_best((uint32_t)(i & 255), i >> 8);
    __syncthreads();
    for (uint32_t g = tid; g < n_pad; g += MIGT_THREADS) {
        uint32_t st = 0;
        if (g < n_gpus) { const kgwe_mig_gpu v = gpus[g]; st = (((v.enabled ? 1u : 0u) | ((v.cls & 1u) << 1)) << 8) | v.occ; }
        state[g] = (uint16_t)st;
        for (int q = 0; q < MIG_GEOS; ++q) {
            const uint32_t b = (st >> 8 & 1u) ? lut[q * 256 + (st & 0xffu)] : 0xFFu;
            leaf[q * n_pad + g] = (uint8_t)((((st >> 9) & 1u) << 7) | (b == 0xFFu ? 0x7Fu : b));
        }
    }
    for (uint32_t i = tid; i < 10 * 32; i += MIGT_THREADS) L2[i] = MIG_KEY_NONE;
    __syncthreads();
    // level 1: tree t = cls * 5 + geometry, node i = min over GPUs 32 i .. 32 i + 31
    for (uint32_t w = warp; w < 5 * n1; w += MIGT_THREADS / 32) {
        const uint32_t t = my_cls * 5u + w / n1, i = w % n1, g = i * 32 + lane;
        const uint32_t k = __reduce_min_sync(0xffffffffu, mig_leaf_key(leaf[(t % 5) * n_pad + g], t / 5, g, state[g]));
        if (lane == 0) L1[t * n1 + i] = k;
    }
    __syncthreads();
    for (uint32_t w = warp; w < 5 * n2; w += MIGT_THREADS / 32) {
        const uint32_t t = my_cls * 5u + w / n2, i = w % n2, j = i * 32 + lane;
        const uint32_t k = __reduce_min_sync(0xffffffffu, j < n1 ? L1[t * n1 + j] : MIG_KEY_NONE);
        if (lane == 0) L2[t * 32 + i] = k;
    }
    __syncthreads();
    for (uint32_t t = my_cls * 5u + warp; t < my_cls * 5u + 5u; t += MIGT_THREADS / 32) {
        const uint32_t k = __reduce_min_sync(0xffffffffu, L2[t * 32 + lane]);
        if (lane == 0) root[t] = k;
    }
    for (uint32_t i = tid; i < 1024u && i < n_reqs; i += MIGT_THREADS) req_stage[i] = profiles[i];
    __syncthreads();
    int profile = n_reqs ? req_stage[0] : 0;
    for (uint32_t r = 0; r < n_reqs; ++r) {
        const int profile_next = req_stage[(r + 1u) & 1023u];      // the next request's profile is fetched now, off the critical path
        const bool known = profile >= 1 && profile <= KGWE_MIG_PROFILES;
        const int geo = known ? mig_geo(profile) : 0, cls = known ? mig_cls(profile) : 0;
        const bool mine = known ? ((uint32_t)cls == my_cls) : (my_cls == 0u);
        const bool refill = (r & 511u) == 511u && r + 513u < n_reqs;
        if (!mine && !refill) { profile = profile_next; continue; }      // the other class's request: nothing of this CTA's changes
        const uint32_t key = (known && mine) ? root[cls * 5 + geo] : MIG_KEY_NONE;
        if (!mine) {
        } else if (key == MIG_KEY_NONE) {                    // nothing fits: nothing changes
            if (tid == 0) { kgwe_mig_assignment a; a.gpu_idx = KGWE_NO_NODE; a.start = a.size = a.cost = 0; a.profile = (uint8_t)profile; out[r] = a; }
        } else {
            const uint32_t g = (key >> 11) & 0xFFFFu, start = key & 7u, size = (uint32_t)mig_size(geo);
            const uint32_t occ_new = ((key >> 3) & 0xffu) | (((1u << size) - 1u) << start);
            const uint32_t b = lut[warp * 256 + occ_new];
            const uint32_t t = (uint32_t)cls * 5u + (uint32_t)warp;
            const uint32_t i1 = g >> 5, i2 = i1 >> 5, gl = (g & ~31u) + lane, j = (i1 & ~31u) + lane;
            const uint32_t leaf_old = leaf[warp * n_pad + gl];
            const uint32_t st_old = state[gl];
            const uint32_t row1 = j < n1 ? L1[t * n1 + j] : MIG_KEY_NONE;
            const uint32_t row2 = L2[t * 32 + lane];
            if (tid == 0) {
                kgwe_mig_assignment a;
                a.gpu_idx = g; a.start = (uint8_t)start; a.size = (uint8_t)size; a.cost = (uint8_t)(key >> 27); a.profile = (uint8_t)profile;
                out[r] = a;
            }
            const uint32_t leaf_new = ((uint32_t)cls << 7) | (b == 0xFFu ? 0x7Fu : b);
            // the three levels of the repaired path: the minima of the rest of each row do not depend on the level below,
            // so the three warp reductions are issued together and only two scalar minima are left on the dependent chain
            const uint32_t rest1 = __reduce_min_sync(0xffffffffu, j == i1 ? MIG_KEY_NONE : row1);
            const uint32_t rest2 = __reduce_min_sync(0xffffffffu, (uint32_t)lane == i2 ? MIG_KEY_NONE : row2);
            const uint32_t k1 = __reduce_min_sync(0xffffffffu, mig_leaf_key(gl == g ? leaf_new : leaf_old, (uint32_t)cls, gl, gl == g ? occ_new : st_old));
            const uint32_t k2 = min(k1, rest1);
            const uint32_t k3 = min(k2, rest2);
            if (lane == 0) { leaf[warp * n_pad + g] = (uint8_t)leaf_new; L1[t * n1 + i1] = k1; L2[t * 32 + i2] = k2; root[t] = k3; if (warp == 0) state[g] = (uint16_t)(0x100u | ((uint32_t)cls << 9) | occ_new); }
        }
        // every 512 requests the half of the ring consumed last is refilled with the profiles 512..1023 requests ahead (the
        // loads are issued by all threads at once; a per-request refill would put a global load in front of every barrier)
        if (refill) {
            const uint32_t first = r + 513u;
            for (uint32_t i = tid; i < 512u && first + i < n_reqs; i += MIGT_THREADS) req_stage[(first + i) & 1023u] = profiles[first + i];
        }
        __syncthreads();                              // warps stay in step; the next request reads the repaired roots
        profile = profile_next;
    }
    for (uint32_t g = tid; g < n_gpus; g += MIGT_THREADS)
        if (((state[g] >> 9) & 1u) == my_cls) gpus[g].occ = (uint8_t)(state[g] & 0xff);      // (this CTA's class: the other CTA never touched these)
}

// Rebalance, device part 1: exclusive rank of every idle GPU (enabled, occ == 0) within its class,
// in index order.  One CTA, chunked ballot/shuffle scan.
__global__ void __launch_bounds__(1024, 1)
mig_rank_kernel(const kgwe_mig_gpu* __restrict__ gpus, uint32_t n_gpus, uint32_t* __restrict__ rank, uint32_t* __restrict__ idle_count) {
    __shared__ uint32_t wsum[2][32];
    __shared__ uint32_t tot[2], carry[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 2) carry[tid] = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_gpus; base += 1024) {
        const uint32_t g = base + tid;
        int c = -1;
        if (g < n_gpus) { const kgwe_mig_gpu v = gpus[g]; if (v.enabled && v.occ == 0) c = v.cls & 1; }
        uint32_t pre[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint32_t b = __ballot_sync(0xffffffffu, c == k);
            pre[k] = __popc(b & ((1u << lane) - 1u));
            if (lane == 0) wsum[k][warp] = __popc(b);
        }
        __syncthreads();
        if (warp == 0) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const uint32_t v = wsum[k][lane]; uint32_t x = v;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, x, off); if (lane >= off) x += o; }
                wsum[k][lane] = x - v;
                if (lane == 31) tot[k] = x;
            }
        }
        __syncthreads();
        if (g < n_gpus) rank[g] = c < 0 ? 0xFFFFFFFFu : carry[c] + wsum[c][warp] + pre[c];
        __syncthreads();
        if (tid < 2) carry[tid] += tot[tid];
        __syncthreads();
    }
    if (tid < 2) idle_count[tid] = carry[tid];
}

// Plan computed on the host from the idle counts and the targets (the CPU checker restates the same rule independently).
struct MigPlan {
    uint32_t first_rank[2][MIG_GEOS + 1];   // class c: ranks [first_rank[c][q], first_rank[c][q+1]) get geometry q
    uint32_t op_base[2][MIG_GEOS];          // op index of the first GPU of (class, geometry)
};

// Rebalance, device part 2: every idle GPU emits the create ops of its profile.
__global__ void __launch_bounds__(256)
mig_emit_kernel(const kgwe_mig_gpu* __restrict__ gpus, uint32_t n_gpus, const uint32_t* __restrict__ rank, MigPlan plan,
                kgwe_mig_op* __restrict__ ops, uint32_t ops_cap) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_gpus) return;
    const uint32_t r = rank[g];
    if (r == 0xFFFFFFFFu) return;
    const int c = gpus[g].cls & 1;
    for (int q = 0; q < MIG_GEOS; ++q) {
        if (r < plan.first_rank[c][q] || r >= plan.first_rank[c][q + 1]) continue;
        const int cnt = mig_nstarts(q);
        const uint32_t at = plan.op_base[c][q] + (r - plan.first_rank[c][q]) * cnt;
        for (int i = 0; i < cnt; ++i)
            if (at + i < ops_cap) {
                kgwe_mig_op op;
                op.gpu_idx = g; op.profile = (uint8_t)(2 * q + 1 + c); op.start = (uint8_t)mig_start(q, i); op.size = (uint8_t)mig_size(q); op.pad = 0;
                ops[at + i] = op;
            }
    }
}

// ---- reconfiguration pass with destroy ops (kgwe_mig_reconfigure) -----------------------------------------
// Placement bit b of the 14 legal placements <-> (geometry, start index).
__host__ __device__ inline int mig_bit_geo(int b) { return b < 7 ? 0 : b < 10 ? 1 : b < 12 ? 2 : b == 12 ? 3 : 4; }
__host__ __device__ inline int mig_bit_first(int geo) { return geo == 0 ? 0 : geo == 1 ? 7 : geo == 2 ? 10 : geo == 3 ? 12 : 13; }
__host__ __device__ inline uint32_t mig_geo_bits(int geo) { return ((1u << mig_nstarts(geo)) - 1u) << mig_bit_first(geo); }

// Part 1: exclusive rank of every reconfigurable GPU (enabled, nothing in use) within its class, in index order.
__global__ void __launch_bounds__(1024, 1)
mig_state_rank_kernel(const kgwe_mig_gpu_state* __restrict__ gpus, uint32_t n_gpus, uint32_t* __restrict__ rank, uint32_t* __restrict__ counts) {
    __shared__ uint32_t wsum[2][32];
    __shared__ uint32_t tot[2], carry[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 2) carry[tid] = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_gpus; base += 1024) {
        const uint32_t g = base + tid;
        int c = -1;
        if (g < n_gpus) { const kgwe_mig_gpu_state v = gpus[g]; if (v.enabled && v.in_use == 0) c = v.cls & 1; }
        uint32_t pre[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint32_t b = __ballot_sync(0xffffffffu, c == k);
            pre[k] = __popc(b & ((1u << lane) - 1u));
            if (lane == 0) wsum[k][warp] = __popc(b);
        }
        __syncthreads();
        if (warp == 0) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const uint32_t v = wsum[k][lane]; uint32_t x = v;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, x, off); if (lane >= off) x += o; }
                wsum[k][lane] = x - v;
                if (lane == 31) tot[k] = x;
            }
        }
        __syncthreads();
        if (g < n_gpus) rank[g] = c < 0 ? 0xFFFFFFFFu : carry[c] + wsum[c][warp] + pre[c];
        __syncthreads();
        if (tid < 2) carry[tid] += tot[tid];
        __syncthreads();
    }
    if (tid < 2) counts[tid] = carry[tid];
}

// Part 2, one launch per (class, geometry) in output order would serialise; instead ONE CTA walks the table once per
// (class, geometry) pair in output order - ten chunked scans over the op counts of the GPUs assigned to that pair - and
// every GPU writes its own ops at the offset the scan gives it.  Also counts the supply.
struct MigRanks { uint32_t first_rank[2][MIG_GEOS + 1]; };
__global__ void __launch_bounds__(1024, 1)
mig_reconf_kernel(const kgwe_mig_gpu_state* __restrict__ gpus, uint32_t n_gpus, const uint32_t* __restrict__ rank, MigRanks plan,
                  kgwe_mig_op* __restrict__ ops, uint32_t ops_cap, uint32_t* __restrict__ n_ops_out, uint32_t* __restrict__ supply_out) {
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry, tot;
    __shared__ uint32_t supply[KGWE_MIG_PROFILES];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    if (tid < KGWE_MIG_PROFILES) supply[tid] = 0;
    __syncthreads();
    // free instances of the GPUs that are left alone
    for (uint32_t g = tid; g < n_gpus; g += 1024) {
        const kgwe_mig_gpu_state v = gpus[g];
        if (!v.enabled || v.in_use == 0) continue;
        uint32_t fr = (uint32_t)v.created & ~(uint32_t)v.in_use & 0x3FFFu;
        while (fr) { const int b = __ffs((int)fr) - 1; fr &= fr - 1u; atomicAdd(&supply[2 * mig_bit_geo(b) + (v.cls & 1)], 1u); }
    }
    for (int c = 0; c < 2; ++c)
        for (int q = 0; q < MIG_GEOS; ++q) {
            const uint32_t r0 = plan.first_rank[c][q], r1 = plan.first_rank[c][q + 1];
            if (r1 == r0) continue;
            if (tid == 0) atomicAdd(&supply[2 * q + c], (r1 - r0) * (uint32_t)mig_nstarts(q));
            const uint32_t want = mig_geo_bits(q);
            for (uint32_t base = 0; base < n_gpus; base += 1024) {
                const uint32_t g = base + tid;
                uint32_t cnt = 0, have = 0;
                bool mine = false;
                if (g < n_gpus) {
                    const uint32_t r = rank[g];
                    const kgwe_mig_gpu_state v = gpus[g];
                    mine = r != 0xFFFFFFFFu && (v.cls & 1) == c && r >= r0 && r < r1;
                    have = (uint32_t)v.created & 0x3FFFu;
                    if (mine) cnt = __popc(have & ~want) + __popc(want & ~have);
                }
                uint32_t x = cnt;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, x, off); if (lane >= off) x += o; }
                if (lane == 31) wsum[warp] = x;
                __syncthreads();
                if (warp == 0) {
                    const uint32_t v = wsum[lane]; uint32_t y = v;
#pragma unroll
                    for (int off = 1; off < 32; off <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, y, off); if (lane >= off) y += o; }
                    wsum[lane] = y - v;
                    if (lane == 31) tot = y;
                }
                __syncthreads();
                if (mine && cnt) {
                    uint32_t at = carry + wsum[warp] + (x - cnt);
                    for (uint32_t d = have & ~want; d; d &= d - 1u, ++at) {
                        const int b = __ffs((int)d) - 1, gq = mig_bit_geo(b);
                        if (at < ops_cap) { kgwe_mig_op op; op.gpu_idx = g; op.profile = (uint8_t)(2 * gq + 1 + c); op.start = (uint8_t)mig_start(gq, b - mig_bit_first(gq)); op.size = (uint8_t)mig_size(gq); op.pad = KGWE_MIG_OP_DESTROY; ops[at] = op; }
                    }
                    for (uint32_t m = want & ~have; m; m &= m - 1u, ++at) {
                        const int b = __ffs((int)m) - 1;
                        if (at < ops_cap) { kgwe_mig_op op; op.gpu_idx = g; op.profile = (uint8_t)(2 * q + 1 + c); op.start = (uint8_t)mig_start(q, b - mig_bit_first(q)); op.size = (uint8_t)mig_size(q); op.pad = KGWE_MIG_OP_CREATE; ops[at] = op; }
                    }
                }
                __syncthreads();
                if (tid == 0) carry += tot;
                __syncthreads();
            }
        }
    __syncthreads();
    if (tid == 0) *n_ops_out = carry;
    if (tid < KGWE_MIG_PROFILES) supply_out[tid] = supply[tid];
}

// GPUs per (class, geometry) from the targets: floor of the shares, the remainder to the largest fractional parts, ties to
// the lowest profile id.  counts[c] GPUs of class c are to be split.
inline void mig_split(const uint32_t counts[2], const double* target, MigRanks& plan) {
    for (int c = 0; c < 2; ++c) {
        double S = 0.0;
        for (int q = 0; q < MIG_GEOS; ++q) { const double t = target[2 * q + c]; if (t > 0.0) S += t; }
        uint32_t n[MIG_GEOS] = {0, 0, 0, 0, 0}; double frac[MIG_GEOS] = {0, 0, 0, 0, 0};
        if (S > 0.0 && counts[c] > 0) {
            uint32_t used = 0;
            for (int q = 0; q < MIG_GEOS; ++q) {
                const double t = target[2 * q + c];
                if (!(t > 0.0)) continue;
                const double share = (t / S) * (double)counts[c];
                n[q] = (uint32_t)share; frac[q] = share - (double)n[q]; used += n[q];
            }
            for (uint32_t left = counts[c] - used; left > 0; --left) {
                int bq = -1;
                for (int q = 0; q < MIG_GEOS; ++q) if (target[2 * q + c] > 0.0 && (bq < 0 || frac[q] > frac[bq])) bq = q;
                if (bq < 0) break;
                n[bq]++; frac[bq] = -1.0;
            }
        }
        uint32_t r = 0;
        for (int q = 0; q < MIG_GEOS; ++q) { plan.first_rank[c][q] = r; r += n[q]; }
        plan.first_rank[c][MIG_GEOS] = r;
    }
}

// Host side of the plan: GPUs per profile from
// the targets, largest-remainder rounding, ties to the lowest profile id.
inline void mig_make_plan(const uint32_t idle[2], const double* target, MigPlan& plan, uint32_t* n_ops, uint32_t* supply) {
    uint32_t ops = 0;
    for (int p = 0; p < KGWE_MIG_PROFILES; ++p) supply[p] = 0;
    for (int c = 0; c < 2; ++c) {
        double S = 0.0;
        for (int q = 0; q < MIG_GEOS; ++q) { const double t = target[2 * q + c]; if (t > 0.0) S += t; }
        uint32_t n[MIG_GEOS] = {0, 0, 0, 0, 0}; double frac[MIG_GEOS] = {0, 0, 0, 0, 0};
        if (S > 0.0 && idle[c] > 0) {
            uint32_t used = 0;
            for (int q = 0; q < MIG_GEOS; ++q) {
                const double t = target[2 * q + c];
                if (!(t > 0.0)) continue;
                const double share = (t / S) * (double)idle[c];
                n[q] = (uint32_t)share; frac[q] = share - (double)n[q]; used += n[q];
            }
            for (uint32_t left = idle[c] - used; left > 0; --left) {
                int bq = -1;
                for (int q = 0; q < MIG_GEOS; ++q) if (target[2 * q + c] > 0.0 && (bq < 0 || frac[q] > frac[bq])) bq = q;
                if (bq < 0) break;
                n[bq]++; frac[bq] = -1.0;
            }
        }
        uint32_t r = 0;
        for (int q = 0; q < MIG_GEOS; ++q) {
            plan.first_rank[c][q] = r; plan.op_base[c][q] = ops;
            r += n[q]; ops += n[q] * (uint32_t)mig_nstarts(q);
            supply[2 * q + c] = n[q] * (uint32_t)mig_nstarts(q);
        }
        plan.first_rank[c][MIG_GEOS] = r;
    }
    *n_ops = ops;
}

}  // namespace kgwe
