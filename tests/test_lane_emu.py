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
    lib.lane_emu_queue_state_check.argtypes = [vp, u32, vp, u32]
    lib.lane_emu_queue_state_check.restype = C.c_longlong
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


def test_queue_kernel_state_refresh(kg, emu):
    """The owner lane of the cluster queue kernel re-derives its node's state after a commit from shared-memory columns,
    without a division (kgwe_queue_state.cuh: the utilisation bonus compares the ordered sum with 30 A and 60 A).  That must
    equal node_ctx_core + lane_node_init on the record for EVERY allocation state: all 256 masks x several counts, on
    clusters that include ragged nodes, unhealthy GPUs, MIG, and utilisations chosen to sit on and next to the 30 / 60
    thresholds of the mean (scheduler.go:540-550) and at the edges of what an upsert accepts."""
    abi = kg.abi
    nodes = np.concatenate([kg.synth.make_cluster(40, "mixed", seed=3), kg.synth.config(1)[0]])
    rng = np.random.default_rng(5)
    edge = nodes[:24].copy()
    for i in range(len(edge)):
        g = int(edge["n_gpus"][i])
        kind = i % 6
        if kind == 0:   # means exactly on the thresholds for some subsets
            edge["util"][i, :g] = rng.choice([30.0, 60.0, 29.999999999999996, 30.000000000000004, 59.99999999999999, 60.00000000000001], g)
        elif kind == 1:  # sums that hit 30 A / 60 A only through rounding
            edge["util"][i, :g] = rng.choice([10.0, 20.0, 40.0, 50.0, 70.0, 80.0, 0.1, 0.2, 59.9, 29.9, 30.1, 60.1], g)
        elif kind == 2:
            edge["util"][i, :g] = rng.choice([-50.0, 0.0, 1e300, 5e-324, 1e-300, 100.0], g)
        elif kind == 3:
            edge["util"][i, :g] = 30.0 + (rng.integers(-3, 4, g) * np.spacing(30.0))
        elif kind == 4:
            edge["util"][i, :g] = 60.0 + (rng.integers(-3, 4, g) * np.spacing(60.0))
        else:
            edge["util"][i, :g] = rng.uniform(0.0, 100.0, g)
        edge["free_bytes"][i, :g] = rng.choice(np.array([0, 1, 2**63, 2**64 - 1, 80 * 2**30], dtype=np.uint64), g)
    nodes = np.ascontiguousarray(np.concatenate([nodes, edge]), dtype=abi.NODE_IMAGE)
    cases = np.array([(m, c) for m in range(256) for c in (0, 1, 3, 7, 8, 9, 200, 2**32 - 1)], dtype=np.uint32)
    bad = emu.lane_emu_queue_state_check(nodes.ctypes.data, len(nodes), cases.ctypes.data, len(cases))
    assert bad == 0
