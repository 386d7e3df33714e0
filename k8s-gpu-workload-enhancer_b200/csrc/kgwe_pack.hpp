// kgwe_pack.hpp — ABI record (kgwe_node_image) <-> device record (DevNode), host side only.
// Used by the engine's upsert path and by the host-side emulation harness of the tests.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstring>

#include "kgwe_device.cuh"

namespace kgwe {

// Uniform-link code of a record (kgwe_device.cuh): set when all its ACTIVE links carry the same bandwidth u > 0
// (whole-number or not).  pad1[0] = 0x80 | m << 3 | c names one such link (c's link to m) so the kernels can fetch u;
// 0x40 = the record has no active link at all (every walk dies at once); 0 = not uniform.
// Derived from the record itself, never taken from outside.
inline void set_uniform_unit(DevNode& rec) {
    rec.pad1[0] = rec.pad1[1] = rec.pad1[2] = 0;
    bool any = false; double u = 0.0; int loc = 0;
    for (int m = 0; m < KGWE_MAX_GPUS; ++m)
        for (int c = 0; c < KGWE_MAX_GPUS; ++c) {
            if (c == m || !(rec.act_in[m] >> c & 1)) continue;
            const double b = rec.bw_in[m][bw_slot(m, c)];
            if (!(b > 0.0) || (any && b != u)) return;
            if (!any) { any = true; u = b; loc = (m << 3) | c; }
        }
    rec.pad1[0] = any ? (uint8_t)(0x80 | loc) : (uint8_t)0x40;
}

// NODE_INTEGRAL of a record, from its own bandwidths (never trusted from outside).
inline bool record_is_integral(const DevNode& rec) {
    for (int m = 0; m < KGWE_MAX_GPUS; ++m)
        for (int c = 0; c < KGWE_MAX_GPUS; ++c) {
            const double b = rec.bw_in[m][c];
            if (b != std::floor(b)) return false;
        }
    return true;
}

// ABI record -> device record: transpose + chunk-swizzle the NVLink matrix (kgwe_device.cuh), transpose the
// active masks, normalise masks to n_gpus.  Returns 0, or -1 with a message in err.
inline int to_device_record(const kgwe_node_image& in, DevNode& out, char* err, size_t err_cap) {
    if (in.n_gpus > KGWE_MAX_GPUS) { snprintf(err, err_cap, "n_gpus %u > %d", in.n_gpus, KGWE_MAX_GPUS); return -1; }
    const int ng = in.n_gpus;
    const uint8_t gm = (uint8_t)((1u << ng) - 1u);
    memset(&out, 0, sizeof out);
    bool integral = true;
    for (int i = 0; i < ng; ++i) {
        const uint8_t present = in.nvl_present[i] & gm & (uint8_t)~(1u << i);
        const uint8_t active = in.nvl_active[i] & present;
        out.present[i] = present;
        for (int j = 0; j < ng; ++j) {
            double bw = (present >> j & 1) ? in.nvl_bw[i][j] : 0.0;
            if (!(bw >= 0.0) || !(bw <= KGWE_MAX_BW_GBPS)) { snprintf(err, err_cap, "nvl_bw[%d][%d]=%g outside [0,%g]", i, j, bw, KGWE_MAX_BW_GBPS); return -1; }
            if (bw == 0.0) bw = 0.0;                              // -0.0 -> +0.0
            if (bw != std::floor(bw)) integral = false;
            out.bw_in[j][bw_slot(j, i)] = bw;                     // i's link to j
            if (active >> j & 1) out.act_in[j] |= (uint8_t)(1u << i);
        }
        const double pb = in.pcie_bw[i], ut = in.util[i];
        if (!(pb >= 0.0) || !(pb <= KGWE_MAX_BW_GBPS)) { snprintf(err, err_cap, "pcie_bw[%d]=%g outside [0,%g]", i, pb, KGWE_MAX_BW_GBPS); return -1; }
        if (!std::isfinite(ut)) { snprintf(err, err_cap, "util[%d] is not finite", i); return -1; }
        out.pcie_bw[i] = pb; out.util[i] = ut; out.free_bytes[i] = in.free_bytes[i];
        out.numa[i] = in.numa[i]; out.pcie_switch[i] = in.pcie_switch[i]; out.arch[i] = in.arch[i];
    }
    memcpy(out.mig_free, in.mig_free, sizeof out.mig_free);
    out.n_gpus = (uint8_t)ng; out.healthy_mask = in.healthy_mask & gm;
    out.mig_enabled_mask = in.mig_enabled_mask & gm; out.mig_has_free_mask = in.mig_has_free_mask & in.mig_enabled_mask & gm;
    out.node_flags = in.node_flags;
    out.flags = NODE_PRESENT | (integral ? NODE_INTEGRAL : 0);
    set_uniform_unit(out);
    return 0;
}

inline void to_abi_record(const DevNode& in, kgwe_node_image& out) {
    memset(&out, 0, sizeof out);
    const int ng = in.n_gpus;
    for (int i = 0; i < ng; ++i) {
        out.nvl_present[i] = in.present[i];
        for (int j = 0; j < ng; ++j) {
            out.nvl_bw[i][j] = in.bw_in[j][bw_slot(j, i)];
            if (in.act_in[j] >> i & 1) out.nvl_active[i] |= (uint8_t)(1u << j);
        }
        out.pcie_bw[i] = in.pcie_bw[i]; out.util[i] = in.util[i]; out.free_bytes[i] = in.free_bytes[i];
        out.numa[i] = in.numa[i]; out.pcie_switch[i] = in.pcie_switch[i]; out.arch[i] = in.arch[i];
    }
    memcpy(out.mig_free, in.mig_free, sizeof out.mig_free);
    out.n_gpus = in.n_gpus; out.healthy_mask = in.healthy_mask; out.allocated_mask = in.allocated_mask;
    out.flags = in.flags & NODE_PRESENT; out.n_allocs = in.n_allocs;
    out.mig_enabled_mask = in.mig_enabled_mask; out.mig_has_free_mask = in.mig_has_free_mask; out.node_flags = in.node_flags;
}

}  // namespace kgwe
