"""K1's per-lane evaluation (csrc/kgwe_lane.cuh) compiled for the host and checked against the oracle.

The scoring kernel runs scalar code per (pod, node) with one lane per node; tests/host/lane_emu.cpp builds that
same code with g++ (stand-ins for the intrinsics) so its logic - every search tier, kernel variant and search mode -
can be compared with the oracle here, without a GPU.  The GPU tests then check the kernel itself."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
MODE_FULL, MODE_PRUNE, MODE_MEMO = 0, 1, 2


@pytest.fixture(scope="module")
def emu():
    subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "host"), "liblane_emu.so"])
    lib = C.CDLL(os.path.join(HERE, "host", "liblane_emu.so"))
    vp, u32 = C.c_void_p, C.c_uint32
    lib.lane_emu_score_batch.argtypes = [vp, vp, u32, vp, u32, vp, u32, vp, vp, C.c_int, C.c_int, u32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    return lib


def run_emu(emu, abi, nodes, ledger, pods, hint, general, tables, mode, cfg=None):
    nodes = np.ascontiguousarray(nodes, dtype=abi.NODE_IMAGE)
    pods = np.ascontiguousarray(pods, dtype=abi.POD_REQ)
    ledger = np.ascontiguousarray(ledger, dtype=abi.ALLOC)
    out = np.zeros((len(pods), len(nodes)), dtype=abi.NODE_SCORE)
    cfg = cfg if cfg is not None else abi.default_config(len(nodes))
    sc, rs = C.c_uint64(0), C.c_uint64(0)
    if hint is not None:
        hint = np.ascontiguousarray(hint, dtype=np.uint8)
    rc = emu.lane_emu_score_batch(C.addressof(cfg), nodes.ctypes.data, len(nodes), ledger.ctypes.data if len(ledger) else None, len(ledger),
                                  pods.ctypes.data, len(pods), hint.ctypes.data if hint is not None else None, out.ctypes.data,
                                  int(general), int(tables), mode, C.byref(sc), C.byref(rs))
    assert rc == 0
    return out, sc.value, rs.value


VARIANTS = [(0, False), (16, False), (32, False), (32, True), (16, True)]      # (matrix units per lane, subset tables)


@pytest.mark.parametrize("num,scale", [(1, 1.0), (2, 0.1), (3, 0.02), (5, 0.02)])
def test_configs(kg, orc, emu, num, scale):
    abi = kg.abi
    nodes, pods, ledger = kg.synth.config(num, scale)
    ref, _, ctr = orc.score_batch(abi, nodes, ledger, pods, want_scores=True, threads=4)
    for general, tables in VARIANTS:
        for mode in (MODE_FULL, MODE_PRUNE, MODE_MEMO):
            got, scored, resolved = run_emu(emu, abi, nodes, ledger, pods, None, general, tables, mode)
            assert got.tobytes() == ref.tobytes(), f"config {num}: general={general} tables={tables} mode={mode}"
            assert resolved == ctr["candidates"]
            if mode == MODE_FULL:
                assert scored == resolved
            else:
                assert scored <= resolved


@pytest.mark.parametrize("seed", range(1, 41))
def test_fuzz_space(kg, orc, emu, seed):
    """The differential fuzzer's input space (tests/fuzz_parity.py): ragged nodes, inactive / zero / fractional /
    irregular links, every preference, architectures, MIG, minimum bandwidth, exhaustive pods, hints, ledgers."""
    import fuzz_parity
    abi = kg.abi
    nodes, pods, led, hint, flags, topk, gang_of, kind = fuzz_parity.make_case(abi, kg.synth, seed, n_lo=20, n_span=120, p_lo=20, p_span=80)
    ref, _, ctr = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=True, threads=4)
    for general, tables in VARIANTS:
        for mode in (MODE_FULL, MODE_PRUNE, MODE_MEMO):
            got, scored, resolved = run_emu(emu, abi, nodes, led, pods, hint, general, tables, mode)
            bad = np.nonzero(got.view(np.uint8).reshape(got.shape + (16,)) != ref.view(np.uint8).reshape(ref.shape + (16,)))
            assert got.tobytes() == ref.tobytes(), (f"seed {seed} (kind {kind}): general={general} tables={tables} mode={mode}: first difference at "
                                                    f"pod {bad[0][0]} node {bad[1][0]}: {got[bad[0][0], bad[1][0]]} vs {ref[bad[0][0], bad[1][0]]}; pod {pods[bad[0][0]]}")
            assert resolved == ctr["candidates"]


def test_non_default_weights(kg, orc, emu):
    abi = kg.abi
    nodes, pods, ledger = kg.synth.config(2, 0.05)
    for wt, wr, wb, bonus, norm in ((0, 0, 0, 10, 900.0), (-40, 35, 25, 10, 450.0), (1000, 1000, 1000, -5, 400.0)):
        cfg = abi.default_config(len(nodes))
        cfg.topology_weight, cfg.resource_weight, cfg.balance_weight, cfg.hint_bonus, cfg.nvlink_norm_gbps = wt, wr, wb, bonus, norm
        hint = kg.synth.make_hint(len(pods), len(nodes), 0.2)
        ref, _, _ = orc.score_batch(abi, nodes, ledger, pods, hint=hint, want_scores=True, cfg=cfg)
        got, _, _ = run_emu(emu, abi, nodes, ledger, pods, hint, 0, False, MODE_FULL, cfg=cfg)
        assert got.tobytes() == ref.tobytes()
