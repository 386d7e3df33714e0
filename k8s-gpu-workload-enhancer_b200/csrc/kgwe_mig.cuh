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
