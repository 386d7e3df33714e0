"""ctypes binding of oracle/libkgwe_oracle.so.  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkgwe_oracle.so")
FAITHFUL, INDEXED = 0, 1
_lib = None


class Counters(C.Structure):
    _fields_ = [("evaluations", C.c_uint64), ("candidates", C.c_uint64), ("seconds", C.c_double)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        vp, u32 = C.c_void_p, C.c_uint32
        _lib.kgwe_oracle_score_batch.argtypes = [vp, vp, u32, vp, u32, vp, u32, vp, vp, u32, vp, C.c_int, C.c_int, C.POINTER(Counters)]
        _lib.kgwe_oracle_topology_hints.argtypes = [vp, vp, u32, vp, u32, vp]
        _lib.kgwe_oracle_topology_gauges.argtypes = [vp, u32, vp]
        _lib.kgwe_oracle_mig_pack.argtypes = [vp, u32, vp, u32, vp]
        _lib.kgwe_oracle_mig_rebalance.argtypes = [vp, u32, vp, vp, u32, C.POINTER(u32), vp]
        _lib.kgwe_oracle_mig_reconfigure.argtypes = [vp, u32, vp, vp, u32, C.POINTER(u32), vp]
        _lib.kgwe_oracle_schedule_gangs.argtypes = [vp, vp, u32, vp, u32, vp, u32, vp, vp, vp, vp, C.POINTER(u32), C.c_int]
        _lib.kgwe_oracle_placement_terms.argtypes = [vp, vp, u32, vp, u32, vp]
        _lib.kgwe_oracle_schedule_queue.argtypes = [vp, vp, u32, vp, u32, vp, u32, vp, vp, vp, C.POINTER(u32), C.c_int, C.POINTER(Counters)]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data


def score_batch(abi, nodes, ledger, pods, hint=None, want_scores=True, topk=0, mode=INDEXED, threads=1, cfg=None):
    """scoreNodes per pod.  Returns (scores[P,N]|None, topk[P,topk]|None, counters dict)."""
    nodes = np.ascontiguousarray(nodes, dtype=abi.NODE_IMAGE)
    pods = np.ascontiguousarray(pods, dtype=abi.POD_REQ)
    ledger = np.ascontiguousarray(ledger if ledger is not None else np.zeros(0, abi.ALLOC), dtype=abi.ALLOC)
    P, N = len(pods), len(nodes)
    scores = np.zeros((P, N), dtype=abi.NODE_SCORE) if want_scores else None
    top = np.zeros((P, topk), dtype=abi.PLACEMENT) if topk else None
    cfg = cfg if cfg is not None else abi.default_config(N)
    if hint is not None:
        hint = np.ascontiguousarray(hint, dtype=np.uint8)
    c = Counters()
    rc = lib().kgwe_oracle_score_batch(C.addressof(cfg), _p(nodes), N, _p(ledger), len(ledger), _p(pods), P, _p(hint),
                                       _p(scores), topk, _p(top), mode, threads, C.byref(c))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return scores, top, {"evaluations": c.evaluations, "candidates": c.candidates, "seconds": c.seconds}


def schedule_queue(abi, nodes, ledger, pods, hint=None, mode=INDEXED, cfg=None):
    """Schedule() per pod in order.  Returns (decisions[P], final ledger, counters dict)."""
    nodes = np.ascontiguousarray(nodes, dtype=abi.NODE_IMAGE)
    pods = np.ascontiguousarray(pods, dtype=abi.POD_REQ)
    ledger = np.ascontiguousarray(ledger if ledger is not None else np.zeros(0, abi.ALLOC), dtype=abi.ALLOC)
    P, N = len(pods), len(nodes)
    out = np.zeros(P, dtype=abi.PLACEMENT)
    led_out = np.zeros(len(ledger) + P, dtype=abi.ALLOC)
    n_out = C.c_uint32(0)
    cfg = cfg if cfg is not None else abi.default_config(N)
    if hint is not None:
        hint = np.ascontiguousarray(hint, dtype=np.uint8)
    c = Counters()
    rc = lib().kgwe_oracle_schedule_queue(C.addressof(cfg), _p(nodes), N, _p(ledger), len(ledger), _p(pods), P, _p(hint),
                                          _p(out), _p(led_out), C.byref(n_out), mode, C.byref(c))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return out, led_out[: n_out.value].copy(), {"evaluations": c.evaluations, "candidates": c.candidates, "seconds": c.seconds}


def schedule_gangs(abi, nodes, ledger, pods, gang_of, hint=None, mode=INDEXED, cfg=None):
    """E5 all-or-nothing gangs.  Returns (decisions[P], final ledger)."""
    nodes = np.ascontiguousarray(nodes, dtype=abi.NODE_IMAGE)
    pods = np.ascontiguousarray(pods, dtype=abi.POD_REQ)
    gang_of = np.ascontiguousarray(gang_of, dtype=np.uint32)
    ledger = np.ascontiguousarray(ledger if ledger is not None else np.zeros(0, abi.ALLOC), dtype=abi.ALLOC)
    P, N = len(pods), len(nodes)
    out = np.zeros(P, dtype=abi.PLACEMENT)
    led_out = np.zeros(len(ledger) + P, dtype=abi.ALLOC)
    n_out = C.c_uint32(0)
    cfg = cfg if cfg is not None else abi.default_config(N)
    if hint is not None:
        hint = np.ascontiguousarray(hint, dtype=np.uint8)
    rc = lib().kgwe_oracle_schedule_gangs(C.addressof(cfg), _p(nodes), N, _p(ledger), len(ledger), _p(pods), P, _p(gang_of), _p(hint),
                                          _p(out), _p(led_out), C.byref(n_out), mode)
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return out, led_out[: n_out.value].copy()


def placement_terms(abi, nodes, placements, cfg=None):
    """E4 side terms by brute force."""
    nodes = np.ascontiguousarray(nodes, dtype=abi.NODE_IMAGE)
    pl = np.ascontiguousarray(placements, dtype=abi.PLACEMENT).reshape(-1)
    out = np.zeros(len(pl), dtype=abi.SIDE_TERMS)
    cfg = cfg if cfg is not None else abi.default_config(len(nodes))
    rc = lib().kgwe_oracle_placement_terms(C.addressof(cfg), _p(nodes), len(nodes), _p(pl), len(pl), _p(out))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return out


def mig_pack(abi, gpus, profiles):
    gpus = np.ascontiguousarray(gpus, dtype=abi.MIG_GPU).copy()
    profiles = np.ascontiguousarray(profiles, dtype=np.uint8)
    out = np.zeros(len(profiles), dtype=abi.MIG_ASSIGNMENT)
    rc = lib().kgwe_oracle_mig_pack(_p(gpus), len(gpus), _p(profiles), len(profiles), _p(out))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return out, gpus


def mig_rebalance(abi, gpus, target, ops_cap=None):
    gpus = np.ascontiguousarray(gpus, dtype=abi.MIG_GPU)
    target = np.ascontiguousarray(target, dtype=np.float64)
    cap = ops_cap if ops_cap is not None else 7 * len(gpus)
    ops = np.zeros(cap, dtype=abi.MIG_OP)
    supply = np.zeros(10, dtype=np.uint32)
    n = C.c_uint32(0)
    rc = lib().kgwe_oracle_mig_rebalance(_p(gpus), len(gpus), _p(target), _p(ops), cap, C.byref(n), _p(supply))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return ops[: min(n.value, cap)].copy(), supply, n.value


def mig_reconfigure(abi, gpus, target, ops_cap=None):
    gpus = np.ascontiguousarray(gpus, dtype=abi.MIG_GPU_STATE)
    target = np.ascontiguousarray(target, dtype=np.float64)
    cap = ops_cap if ops_cap is not None else 21 * len(gpus)
    ops = np.zeros(cap, dtype=abi.MIG_OP)
    supply = np.zeros(10, dtype=np.uint32)
    n = C.c_uint32(0)
    rc = lib().kgwe_oracle_mig_reconfigure(_p(gpus), len(gpus), _p(target), _p(ops), cap, C.byref(n), _p(supply))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return ops[: min(n.value, cap)].copy(), supply, n.value


def topology_hints(abi, nodes, reqs, cfg=None):
    nodes = np.ascontiguousarray(nodes, dtype=abi.NODE_IMAGE)
    reqs = np.ascontiguousarray(reqs, dtype=abi.HINT_REQ)
    out = np.zeros(len(reqs), dtype=abi.TOPOLOGY_HINT)
    cfg = cfg if cfg is not None else abi.default_config(max(1, len(nodes)))
    rc = lib().kgwe_oracle_topology_hints(C.addressof(cfg), _p(nodes), len(nodes), _p(reqs), len(reqs), _p(out))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return out


def topology_gauges(abi, nodes):
    nodes = np.ascontiguousarray(nodes, dtype=abi.NODE_IMAGE)
    out = np.zeros(len(nodes), dtype=np.int32)
    rc = lib().kgwe_oracle_topology_gauges(_p(nodes), len(nodes), _p(out))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return out
