"""Binary layouts of include/kgwe_engine.h as numpy dtypes, and the ctypes loader of the
CUDA engine library.  Plumbing only: no scoring logic lives in Python.

The product path has no CPU fallback: `load_engine_lib()` raises if
csrc/libkgwe_engine.so is missing, and `Engine(...)` raises if the library cannot get a
CUDA device (kgwe_engine_create -> KGWE_ERR_NO_DEVICE).
"""
import ctypes as C
import os

import numpy as np

MAX_GPUS = 8
MIG_PROFILES = 10
MAX_TOPK = 16
NO_NODE = 0xFFFFFFFF

TOPO_NONE, TOPO_NVLINK_OPTIMAL, TOPO_NVLINK_REQUIRED, TOPO_SAME_NUMA, TOPO_SAME_PCIE_SWITCH = range(5)
ARCH_ANY, ARCH_VOLTA, ARCH_TURING, ARCH_AMPERE, ARCH_HOPPER, ARCH_ADA, ARCH_BLACKWELL = range(7)
MIG_NONE, MIG_UNKNOWN = 0, 255
MIG_NAMES = ["1g.5gb", "1g.10gb", "2g.10gb", "2g.20gb", "3g.20gb", "3g.40gb", "4g.20gb", "4g.40gb", "7g.40gb", "7g.80gb"]
POD_EXHAUSTIVE = 0x01
CFG_SUBSET_TABLES = 0x1
CFG_MEMOISE_GROUPS = 0x2
CFG_PRUNE_TIED_STARTS = 0x4
F_ELIGIBLE, F_KEPT, F_HINT, F_FALLBACK, F_MINBW, F_GANG_ROLLBACK = 0x01, 0x02, 0x04, 0x08, 0x10, 0x20
MIG_GPU = np.dtype([("occ", "u1"), ("cls", "u1"), ("enabled", "u1"), ("pad", "u1")])
MIG_ASSIGNMENT = np.dtype([("gpu_idx", "<u4"), ("start", "u1"), ("size", "u1"), ("cost", "u1"), ("profile", "u1")])
MIG_OP = np.dtype([("gpu_idx", "<u4"), ("profile", "u1"), ("start", "u1"), ("size", "u1"), ("pad", "u1")])      # pad: MIG_OP_CREATE / MIG_OP_DESTROY
MIG_GPU_STATE = np.dtype([("created", "<u2"), ("in_use", "<u2"), ("cls", "u1"), ("enabled", "u1"), ("pad", "u1", (2,))])
MIG_OP_CREATE, MIG_OP_DESTROY = 0, 1
SIDE_TERMS = np.dtype([("bisection_gbps", "<f8"), ("ring_bottleneck_gbps", "<f8")])

NODE_IMAGE = np.dtype([
    ("nvl_bw", "<f8", (MAX_GPUS, MAX_GPUS)), ("pcie_bw", "<f8", (MAX_GPUS,)),
    ("free_bytes", "<u8", (MAX_GPUS,)), ("util", "<f8", (MAX_GPUS,)),
    ("nvl_present", "u1", (MAX_GPUS,)), ("nvl_active", "u1", (MAX_GPUS,)),
    ("numa", "i1", (MAX_GPUS,)), ("pcie_switch", "u1", (MAX_GPUS,)), ("arch", "u1", (MAX_GPUS,)),
    ("mig_free", "u1", (MIG_PROFILES,)), ("n_gpus", "u1"), ("healthy_mask", "u1"),
    ("allocated_mask", "u1"), ("flags", "u1"), ("mig_enabled_mask", "u1"), ("mig_has_free_mask", "u1"), ("n_allocs", "<u4"),
    ("node_flags", "u1"), ("pad1", "u1", (3,)),
])
NODE_NVSWITCH, NODE_ANY_NVLINK = 0x01, 0x02
HINT_REQ = np.dtype([("count", "<i4"), ("min_memory_gb", "<i4"), ("prefer_nvlink", "u1"), ("prefer_same_numa", "u1"), ("arch", "u1"), ("pad", "u1", (5,))])
TOPOLOGY_HINT = np.dtype([("node_idx", "<u4"), ("score", "<i4"), ("order", "<u4"), ("gpu_mask", "u1"), ("pad", "u1", (3,)), ("bandwidth_estimate", "<f8")])
POD_REQ = np.dtype([
    ("wl_id", "<u8"), ("gpu_count", "<i4"), ("min_memory_gb", "<i4"), ("min_bandwidth_gbps", "<f8"),
    ("topology_pref", "u1"), ("arch", "u1"), ("mig_profile", "u1"), ("flags", "u1"), ("mig_count", "<i4"),
])
NODE_SCORE = np.dtype([
    ("total", "<i4"), ("topology", "<i2"), ("resource", "<i2"), ("balance", "<i2"),
    ("gpu_mask", "u1"), ("flags", "u1"), ("order", "<u4"),
])
PLACEMENT = np.dtype([
    ("node_idx", "<u4"), ("total", "<i4"), ("topology", "<i2"), ("resource", "<i2"), ("balance", "<i2"),
    ("gpu_mask", "u1"), ("flags", "u1"), ("order", "<u4"), ("reserved", "<u4"), ("est_bandwidth_gbps", "<f8"),
])
ALLOC = np.dtype([("wl_id", "<u8"), ("node_idx", "<u4"), ("gpu_mask", "u1"), ("pad", "u1", (3,))])

assert NODE_IMAGE.itemsize == 768 and POD_REQ.itemsize == 32 and NODE_SCORE.itemsize == 16
assert PLACEMENT.itemsize == 32 and ALLOC.itemsize == 16 and HINT_REQ.itemsize == 16 and TOPOLOGY_HINT.itemsize == 24


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("max_nodes", C.c_uint32),
                ("node_base", C.c_uint32), ("topology_weight", C.c_int32), ("resource_weight", C.c_int32),
                ("balance_weight", C.c_int32), ("hint_bonus", C.c_int32), ("nvlink_norm_gbps", C.c_double),
                ("flags", C.c_uint32), ("reserved", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("evaluations", C.c_uint64), ("candidates", C.c_uint64), ("scheduled", C.c_uint64),
                ("failed", C.c_uint64), ("topology_optimal", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("last_call_gpu_ms", C.c_double),
                ("n_nodes", C.c_uint32), ("n_allocations", C.c_uint32),
                ("score_kernel_ms", C.c_double), ("score_kernel_timed", C.c_uint64), ("candidates_resolved", C.c_uint64),
                ("latency_us_bucket", C.c_uint64 * 10), ("latency_count", C.c_uint64), ("latency_sum_us", C.c_double)]


LATENCY_BUCKET_EDGES_US = [10, 25, 50, 100, 250, 500, 1000, 2500, 5000, float("inf")]


def default_config(max_nodes, device=0, node_base=0):
    """DefaultSchedulerConfig() (reference src/scheduler/types.go:379-392): weights 40/35/25."""
    return Config(C.sizeof(Config), device, max_nodes, node_base, 40, 35, 25, 10, 900.0, 0, 0)


# Every symbol include/kgwe_engine.h declares; tests check the library exports all of them.
ENGINE_SYMBOLS = [
    "kgwe_engine_create", "kgwe_engine_destroy", "kgwe_config_default", "kgwe_node_upsert", "kgwe_nodes_upsert",
    "kgwe_node_remove", "kgwe_alloc_commit", "kgwe_alloc_release", "kgwe_score_batch", "kgwe_score_batch_device",
    "kgwe_topk_merge_device", "kgwe_schedule_queue", "kgwe_node_read", "kgwe_get_stats", "kgwe_last_error",
    "kgwe_version", "kgwe_score_top1_noalloc", "kgwe_topk_merge_host",
    "kgwe_schedule_gangs", "kgwe_placement_terms", "kgwe_mig_pack", "kgwe_mig_rebalance",
    "kgwe_topology_hint_batch", "kgwe_topology_gauges", "kgwe_score_batch_device_p2p",
    "kgwe_snapshot_size", "kgwe_snapshot_save", "kgwe_snapshot_load", "kgwe_node_upsert_remap", "kgwe_score_merge_p2p", "kgwe_mig_reconfigure",
    "kgwe_schedule_queue_p2p",
]
NOALLOC_MAX_PODS = 16384

_HERE = os.path.dirname(os.path.abspath(__file__))
ENGINE_LIB_PATH = os.path.join(_HERE, "csrc", "libkgwe_engine.so")
_lib = None


def load_engine_lib():
    """dlopen csrc/libkgwe_engine.so (built by __graft_entry__.build()).  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(ENGINE_LIB_PATH):
        raise RuntimeError(f"CUDA engine library missing: {ENGINE_LIB_PATH}; run __graft_entry__.build() — there is no CPU fallback")
    lib = C.CDLL(ENGINE_LIB_PATH)
    vp, u32, u64, u8 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint8
    lib.kgwe_engine_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    lib.kgwe_engine_destroy.argtypes = [vp]
    lib.kgwe_config_default.argtypes = [C.POINTER(Config)]
    lib.kgwe_config_default.restype = None
    lib.kgwe_node_upsert.argtypes = [vp, u32, vp]
    lib.kgwe_nodes_upsert.argtypes = [vp, u32, u32, vp]
    lib.kgwe_node_upsert_remap.argtypes = [vp, u32, vp, vp]
    lib.kgwe_node_remove.argtypes = [vp, u32]
    lib.kgwe_alloc_commit.argtypes = [vp, u64, u32, u8]
    lib.kgwe_alloc_release.argtypes = [vp, u64]
    lib.kgwe_score_batch.argtypes = [vp, vp, u32, vp, vp, u32, vp]
    lib.kgwe_score_batch_device.argtypes = [vp, vp, u32, vp, vp, u32, vp, vp]
    lib.kgwe_score_top1_noalloc.argtypes = [vp, vp, u32, vp, vp, vp, vp]
    lib.kgwe_snapshot_size.argtypes = [vp, C.POINTER(u64)]
    lib.kgwe_snapshot_save.argtypes = [vp, vp, u64, C.POINTER(u64)]
    lib.kgwe_snapshot_load.argtypes = [vp, vp, u64]
    lib.kgwe_score_batch_device_p2p.argtypes = [vp, vp, u32, vp, vp, vp, u32, u32, vp]
    lib.kgwe_score_merge_p2p.argtypes = [vp, vp, u32, vp, vp, vp, vp, u32, u32, u32, vp, vp]
    lib.kgwe_topology_hint_batch.argtypes = [vp, vp, u32, vp]
    lib.kgwe_topology_gauges.argtypes = [vp, vp, u32]
    lib.kgwe_mig_pack.argtypes = [vp, vp, u32, vp, u32, vp]
    lib.kgwe_mig_rebalance.argtypes = [vp, vp, u32, vp, vp, u32, C.POINTER(u32), vp]
    lib.kgwe_mig_reconfigure.argtypes = [vp, vp, u32, vp, vp, u32, C.POINTER(u32), vp]
    lib.kgwe_schedule_gangs.argtypes = [vp, vp, u32, vp, vp, vp]
    lib.kgwe_placement_terms.argtypes = [vp, vp, u32, vp]
    lib.kgwe_topk_merge_host.argtypes = [vp, u32, u32, u32, vp]
    lib.kgwe_topk_merge_device.argtypes = [vp, vp, u32, u32, u32, vp, vp]
    lib.kgwe_schedule_queue.argtypes = [vp, vp, u32, vp, vp]
    lib.kgwe_schedule_queue_p2p.argtypes = [vp, vp, u32, vp, vp, u32, u32, u32, vp]
    lib.kgwe_node_read.argtypes = [vp, u32, vp]
    lib.kgwe_get_stats.argtypes = [vp, C.POINTER(Stats)]
    lib.kgwe_last_error.argtypes = [vp]
    lib.kgwe_last_error.restype = C.c_char_p
    lib.kgwe_version.restype = C.c_char_p
    for name in ENGINE_SYMBOLS:
        fn = getattr(lib, name)
        if fn.restype is C.c_int:
            fn.restype = C.c_int
    _lib = lib
    return lib


def ptr(a):
    """Host address of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data
