"""Parity tests proper: the CUDA engine, called through the C ABI, against the CPU oracle
on the same inputs.  Bar: every integer/byte field bit-exact; EstimatedBandwidth (the only
float output) must also be bit-exact here, tolerance 1e-6 relative is the stated ceiling."""
import json
import os

import numpy as np
import pytest

from helpers import CUBE_EDGES, ledger_of, make_node, make_pod, order_list
from test_oracle_kats import KATS, _node, check_expect

pytestmark = pytest.mark.gpu
REL_TOL = 1e-6


def assert_same_placements(a, b):
    for f in a.dtype.names:
        if f == "est_bandwidth_gbps":
            np.testing.assert_allclose(a[f], b[f], rtol=REL_TOL, atol=0)
        else:
            assert np.array_equal(a[f], b[f]), f
    assert a.tobytes() == b.tobytes()      # and in fact bit-identical


def engine_for(kg, nodes, ledger=None, **kw):
    eng = kg.Engine(max_nodes=max(1, len(nodes)), **kw)
    if len(nodes):
        eng.upsert_nodes(nodes)
    if ledger is not None:
        eng.seed_ledger(ledger)
    return eng


@pytest.mark.parametrize("case", KATS["cases"], ids=[c["name"].split(" ")[0] + f"-{i}" for i, c in enumerate(KATS["cases"])])
def test_kat_through_engine(kg, case):
    abi = kg.abi
    nodes = np.array([_node(abi, case["node"])], dtype=abi.NODE_IMAGE)
    pods = np.array([make_pod(abi, **case["pod"])], dtype=abi.POD_REQ)
    eng = engine_for(kg, nodes, ledger_of(abi, [tuple(x) for x in case["ledger"]]))
    scores, top = eng.score_batch(pods, 1, want_scores=True, topk=1)
    check_expect(abi, scores[0, 0], top[0, 0], case["expect"])
    eng.close()


@pytest.mark.parametrize("num,scale,npods", [(1, 1.0, 100), (2, 0.2, 200), (3, 0.03, 300), (5, 0.02, 160)])
@pytest.mark.parametrize("with_hint", [False, True])
def test_stateless_matrix_and_topk(kg, orc, num, scale, npods, with_hint):
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(num, scale)
    pods = pods[:npods]
    hint = synth.make_hint(len(pods), len(nodes), 0.03) if with_hint else None
    eng = engine_for(kg, nodes, led, flags=abi.CFG_MEMOISE_GROUPS if with_hint else 0)     # both group-search modes
    for topk in (1, 5, 16):
        scores, top = eng.score_batch(pods, len(nodes), hint=hint, want_scores=True, topk=topk)
        o_scores, o_top, ctr = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=True, topk=topk, threads=4)
        assert scores.tobytes() == o_scores.tobytes()
        assert_same_placements(top, o_top)
    # top-k without the matrix (fused arg-max path for k=1, internal scratch for k>1)
    for topk in (1, 3):
        _, top = eng.score_batch(pods, len(nodes), hint=hint, want_scores=False, topk=topk)
        _, o_top, _ = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=False, topk=topk, threads=4)
        assert_same_placements(top, o_top)
    st = eng.stats()
    assert st["evaluations"] == 5 * len(pods) * len(nodes)
    assert st["candidates_resolved"] == 5 * ctr["candidates"]          # the reference's counting rule, whatever the search mode
    if not with_hint:
        assert st["candidates"] == 5 * ctr["candidates"]               # full search: every candidate's score was computed
    else:
        assert st["candidates"] < st["candidates_resolved"]            # memoised winners: looked up, not scored again
    eng.close()


@pytest.mark.parametrize("num,scale,npods", [(1, 1.0, 100), (2, 0.1, 150), (3, 0.02, 400), (5, 0.01, 96)])
def test_stateful_queue(kg, orc, num, scale, npods):
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(num, scale)
    pods = pods[:npods]
    eng = engine_for(kg, nodes, led)
    dec = eng.schedule_queue(pods)
    o_dec, o_led, _ = orc.schedule_queue(abi, nodes, led, pods)
    assert_same_placements(dec, o_dec)
    # the device image now carries the oracle's final ledger projection
    alloc = np.zeros(len(nodes), np.uint8)
    cnt = np.zeros(len(nodes), np.uint32)
    for a in o_led:
        alloc[a["node_idx"]] |= a["gpu_mask"]
        cnt[a["node_idx"]] += 1
    for n in range(len(nodes)):
        rec = eng.read_node(n)
        assert rec["allocated_mask"] == alloc[n] and rec["n_allocs"] == cnt[n]
    st = eng.stats()
    assert st["scheduled"] == int((o_dec["node_idx"] != abi.NO_NODE).sum())
    assert st["failed"] == int((o_dec["node_idx"] == abi.NO_NODE).sum())
    assert st["n_allocations"] == len(o_led)
    # scoring after the queue sees the commits
    s2, _ = eng.score_batch(pods[:20], len(nodes), want_scores=True)
    o2, _, _ = orc.score_batch(abi, nodes, o_led, pods[:20], want_scores=True)
    assert s2.tobytes() == o2.tobytes()
    eng.close()


def test_stateful_with_hint_and_multi_node_groups(kg, orc):
    """More nodes than resident lane groups forces the multi-node-per-group path of the queue kernel."""
    abi, synth = kg.abi, kg.synth
    nodes = synth.make_cluster(24000, "mixed", seed=99)
    pods = synth.make_pods(40, seed=99)
    led = synth.make_ledger(nodes, 0.2, seed=99)
    hint = synth.make_hint(len(pods), len(nodes), 0.001, seed=99)
    eng = engine_for(kg, nodes, led)
    dec = eng.schedule_queue(pods, hint=hint)
    o_dec, _, _ = orc.schedule_queue(abi, nodes, led, pods, hint=hint)
    assert_same_placements(dec, o_dec)
    eng.close()


def test_exhaustive_extension(kg, orc):
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(3, 0.02)
    pods = pods[:200].copy()
    eng = engine_for(kg, nodes, led)
    g_scores, _ = eng.score_batch(pods, len(nodes), want_scores=True)
    pods["flags"] = abi.POD_EXHAUSTIVE
    scores, top = eng.score_batch(pods, len(nodes), want_scores=True, topk=2)
    o_scores, o_top, _ = orc.score_batch(abi, nodes, led, pods, want_scores=True, topk=2, threads=4)
    assert scores.tobytes() == o_scores.tobytes()
    assert_same_placements(top, o_top)
    # property: on symmetric bandwidth matrices the exhaustive topology score is >= the greedy one
    assert (scores["topology"] >= g_scores["topology"]).all()
    # greedy and exhaustive pods interleaved inside the same 8-pod octets, a few nodes with fractional
    # bandwidths (generic FP64 path), through the host API and through the device API of an engine
    # created with KGWE_CFG_SUBSET_TABLES - and of one without it (slow path, same answers)
    import torch
    mixed = pods.copy()
    mixed["flags"] = np.where(np.arange(len(mixed)) % 3 == 0, 0, abi.POD_EXHAUSTIVE).astype(np.uint8)
    mixed["topology_pref"][::5] = abi.TOPO_SAME_NUMA
    nodes2 = nodes.copy()
    nodes2["nvl_bw"][::7] *= 1.03125
    o_scores, o_top, _ = orc.score_batch(abi, nodes2, led, mixed, want_scores=True, topk=1, threads=4)
    for flags in (abi.CFG_SUBSET_TABLES, 0, abi.CFG_MEMOISE_GROUPS, abi.CFG_MEMOISE_GROUPS | abi.CFG_SUBSET_TABLES):
        e2 = engine_for(kg, nodes2, led, flags=flags)
        s2, t2 = e2.score_batch(mixed, len(nodes2), want_scores=True, topk=1)
        assert s2.tobytes() == o_scores.tobytes()
        assert_same_placements(t2, o_top)
        d_pods = torch.from_numpy(mixed.view(np.uint8).reshape(len(mixed), abi.POD_REQ.itemsize)).cuda()
        d_sc = torch.zeros((len(mixed), len(nodes2), abi.NODE_SCORE.itemsize), dtype=torch.uint8, device="cuda")
        d_top = torch.zeros((len(mixed), abi.PLACEMENT.itemsize), dtype=torch.uint8, device="cuda")
        e2.score_batch_device(d_pods.data_ptr(), len(mixed), None, d_sc.data_ptr(), 1, d_top.data_ptr())
        torch.cuda.synchronize()
        assert d_sc.cpu().numpy().tobytes() == o_scores.tobytes()
        assert d_top.cpu().numpy().tobytes() == o_top.tobytes()
        e2.close()
    eng.close()


def test_edge_cases(kg, orc):
    abi = kg.abi
    # empty cluster, empty pod list
    eng = kg.Engine(max_nodes=4)
    pods = np.array([make_pod(abi, 2)], dtype=abi.POD_REQ)
    _, top = eng.score_batch(pods, 0, want_scores=False, topk=2)
    assert (top["node_idx"] == abi.NO_NODE).all()
    dec = eng.schedule_queue(pods)
    assert dec[0]["node_idx"] == abi.NO_NODE
    s, t = eng.score_batch(pods[:0], 0, want_scores=True, topk=1)
    assert s.shape == (0, 0)
    # the device entry point on an empty image writes "no node" placements too (not stale memory)
    import torch
    d_pods = torch.from_numpy(pods.view(np.uint8).reshape(1, abi.POD_REQ.itemsize)).cuda()
    for k in (1, 3):
        d_top = torch.full((k, abi.PLACEMENT.itemsize), 0x5A, dtype=torch.uint8, device="cuda")
        eng.score_batch_device(d_pods.data_ptr(), 1, None, None, k, d_top.data_ptr())
        torch.cuda.synchronize()
        got = d_top.cpu().numpy().view(abi.PLACEMENT).reshape(-1)
        assert (got["node_idx"] == abi.NO_NODE).all() and (got["total"] == 0).all() and (got["gpu_mask"] == 0).all()
    # ragged nodes: 0..8 GPUs, all-unhealthy, no links
    nodes = np.array([make_node(abi, n_gpus=g) for g in range(9)] + [make_node(abi, unhealthy=range(8)), make_node(abi, links="none")],
                     dtype=abi.NODE_IMAGE)
    eng2 = engine_for(kg, nodes)
    pods = np.array([make_pod(abi, k, pref=p, mem=m) for k in (0, 1, 2, 5, 8, 9, 100) for p in ("None", "NVLinkOptimal", "SameNUMA", "SamePCIeSwitch")
                     for m in (0, 16, 2000)], dtype=abi.POD_REQ)
    scores, top = eng2.score_batch(pods, len(nodes), want_scores=True, topk=16)
    o_scores, o_top, _ = orc.score_batch(abi, nodes, None, pods, want_scores=True, topk=16)
    assert scores.tobytes() == o_scores.tobytes()
    assert_same_placements(top, o_top)
    # invalid arguments are rejected, not computed
    bad = pods[:1].copy(); bad["gpu_count"] = -1
    with pytest.raises(kg.EngineError):
        eng2.score_batch(bad, len(nodes))
    nan_node = make_node(abi); nan_node["nvl_bw"][0][1] = np.nan
    with pytest.raises(kg.EngineError):
        eng2.upsert_node(0, nan_node)
    with pytest.raises(kg.EngineError):
        eng2.upsert_node(10_000, make_node(abi))
    eng.close(); eng2.close()


def test_incremental_updates_and_ledger(kg, orc):
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(2, 0.05)
    pods = pods[:50]
    eng = engine_for(kg, nodes, led)
    # informer event: one node changes (a GPU goes unhealthy, memory drops)
    nodes = nodes.copy()
    nodes[7]["healthy_mask"] &= 0xF0
    nodes[7]["free_bytes"][:] = 1 << 30
    eng.upsert_node(7, nodes[7])
    # node 3 deleted (discovery.go:597-600): the oracle sees it as an absent record
    eng.remove_node(3)
    o_nodes = nodes.copy(); o_nodes[3]["n_gpus"] = 0; o_nodes[3]["healthy_mask"] = 0
    s, _ = eng.score_batch(pods, len(nodes), want_scores=True)
    o, _, _ = orc.score_batch(abi, o_nodes, led, pods, want_scores=True)
    keep = np.arange(len(nodes)) != 3
    assert s[:, keep].tobytes() == o[:, keep].tobytes()
    assert (s[:, 3]["flags"] == 0).all()
    # node 3 comes back: the ledger entries that name it match again (name-keyed ledger, scheduler.go:587-593)
    eng.upsert_node(3, nodes[3])
    s_back, _ = eng.score_batch(pods, len(nodes), want_scores=True)
    o_back, _, _ = orc.score_batch(abi, nodes, led, pods, want_scores=True)
    assert s_back.tobytes() == o_back.tobytes()
    eng.remove_node(3)
    assert eng.stats()["n_nodes"] == len(nodes)
    # ledger: release / unknown release / same wl_id overwrites (scheduler.go:699, 710-716)
    first = led[0]
    eng.release(first["wl_id"])
    with pytest.raises(kg.EngineError) as ei:
        eng.release(first["wl_id"])
    assert ei.value.code == -5 and "allocation not found" in str(ei.value)
    led2 = led[1:].copy()
    taken5 = int(np.bitwise_or.reduce(led2["gpu_mask"][1:][led2["node_idx"][1:] == 5], initial=0))
    free5 = [g for g in range(int(nodes[5]["n_gpus"])) if not taken5 >> g & 1]
    move = (1 << free5[0]) | (1 << free5[-1])
    if taken5:
        with pytest.raises(kg.EngineError) as ei:              # a GPU another workload holds: refused (scheduler.go:634-640)
            eng.commit(led2[0]["wl_id"], 5, move | (taken5 & -taken5))
        assert ei.value.code == -6
    eng.commit(led2[0]["wl_id"], 5, move)                      # overwrite: moves to node 5
    led2[0]["node_idx"], led2[0]["gpu_mask"] = 5, move
    s, _ = eng.score_batch(pods, len(nodes), want_scores=True)
    o, _, _ = orc.score_batch(abi, o_nodes, led2, pods, want_scores=True)
    assert s[:, keep].tobytes() == o[:, keep].tobytes()
    # a queue that re-schedules a workload already in the ledger
    again = pods[:6].copy()
    again["wl_id"][2] = led2[3]["wl_id"]
    again["wl_id"][4] = again["wl_id"][1]
    dec = eng.schedule_queue(again)
    o_dec, o_led, _ = orc.schedule_queue(abi, o_nodes, led2, again)
    assert_same_placements(dec, o_dec)
    assert eng.stats()["n_allocations"] == len(o_led)
    s, _ = eng.score_batch(pods, len(nodes), want_scores=True)
    o, _, _ = orc.score_batch(abi, o_nodes, o_led, pods, want_scores=True)
    assert s[:, keep].tobytes() == o[:, keep].tobytes()
    # a repeated id whose FIRST occurrence finds no node (9 GPUs asked): nothing to overwrite when the second lands
    twice = pods[10:14].copy()
    twice["wl_id"][3] = twice["wl_id"][0]
    twice["gpu_count"][0] = 9
    dec = eng.schedule_queue(twice)
    o_dec2, o_led2, _ = orc.schedule_queue(abi, o_nodes, o_led, twice)
    assert_same_placements(dec, o_dec2)
    assert dec[0]["node_idx"] == abi.NO_NODE and dec[3]["node_idx"] != abi.NO_NODE
    assert eng.stats()["n_allocations"] == len(o_led2)
    eng.close()


def test_node_read_roundtrip(kg):
    abi, synth = kg.abi, kg.synth
    nodes = synth.make_cluster(40, "mixed", seed=5, partial_node_frac=0.3, mig_frac=0.5)
    eng = engine_for(kg, nodes)
    for n in (0, 13, 39):
        rec = eng.read_node(n)
        for f in ("pcie_bw", "free_bytes", "util", "nvl_present", "nvl_active", "numa", "pcie_switch", "arch", "mig_free", "n_gpus", "healthy_mask"):
            assert np.array_equal(rec[f], nodes[n][f]), f
        assert np.array_equal(rec["nvl_bw"], nodes[n]["nvl_bw"])
    eng.close()


def test_mig_and_arch_filters_synthetic(kg, orc):
    abi, synth = kg.abi, kg.synth
    nodes = synth.make_cluster(300, "mixed", seed=11, partial_node_frac=0.2, mig_frac=0.6, unhealthy_frac=0.1, inactive_link_frac=0.05)
    pods = synth.make_pods(300, seed=11, arch_frac=0.5, mig_frac=0.5, minbw_choices=(0.0, 300.0, 700.0, 1000.0))
    led = synth.make_ledger(nodes, 0.5, seed=11)
    eng = engine_for(kg, nodes, led)
    s, top = eng.score_batch(pods, len(nodes), want_scores=True, topk=8)
    o, o_top, _ = orc.score_batch(abi, nodes, led, pods, want_scores=True, topk=8, threads=4)
    assert s.tobytes() == o.tobytes()
    assert_same_placements(top, o_top)
    assert (s["flags"] & abi.F_MINBW).any() and (s["flags"] & abi.F_FALLBACK).any()
    dec = eng.schedule_queue(pods)
    o_dec, _, _ = orc.schedule_queue(abi, nodes, led, pods)
    assert_same_placements(dec, o_dec)
    eng.close()


def test_full_size_properties(kg, orc):
    """BASELINE config 3 at full size (10k nodes x 10k pods): oracle spot-checks plus
    size-independent properties (arg-max of the matrix row == fused top-1, idempotence,
    no GPU handed out twice by the stateful queue)."""
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(3)
    eng = engine_for(kg, nodes, led)
    N = len(nodes)
    _, top_a = eng.score_batch(pods, N, want_scores=False, topk=1)
    _, top_b = eng.score_batch(pods, N, want_scores=False, topk=1)
    assert top_a.tobytes() == top_b.tobytes()
    sample = np.arange(0, len(pods), 97)
    s, top = eng.score_batch(pods[sample], N, want_scores=True, topk=1)
    o, o_top, _ = orc.score_batch(abi, nodes, led, pods[sample], want_scores=True, topk=1, threads=8)
    assert s.tobytes() == o.tobytes()
    assert_same_placements(top, o_top)
    assert top.tobytes() == top_a[sample].tobytes()
    kept = (s["flags"] & abi.F_KEPT) != 0
    best = np.where(kept, s["total"], -1).argmax(axis=1)        # first max = lowest index
    assert np.array_equal(best.astype(np.uint32), top[:, 0]["node_idx"])
    dec = eng.schedule_queue(pods[:3000])
    o_dec, _, _ = orc.schedule_queue(abi, nodes, led, pods[:3000])
    assert_same_placements(dec, o_dec)
    used = {}
    for a in led:
        used[int(a["node_idx"])] = used.get(int(a["node_idx"]), 0) | int(a["gpu_mask"])
    for d in dec:
        if d["node_idx"] == abi.NO_NODE:
            continue
        n = int(d["node_idx"])
        assert used.get(n, 0) & int(d["gpu_mask"]) == 0
        used[n] = used.get(n, 0) | int(d["gpu_mask"])
    eng.close()


def test_fractional_bandwidths_take_the_fp64_path(kg, orc):
    """Whole-number bandwidths use the integer chains; anything else must fall back to the
    FP64 chains, also when both kinds of node share a warp.  Random doubles make the order of
    the float64 additions matter, so this also pins the summation order."""
    abi, synth = kg.abi, kg.synth
    nodes = synth.make_cluster(700, "mixed", seed=21, unhealthy_frac=0.05, inactive_link_frac=0.03)
    rng = synth.SplitMix64(77)
    frac = rng.uniform(len(nodes)) < 0.5
    jitter = 0.5 + rng.uniform(len(nodes) * 64).reshape(len(nodes), 8, 8)     # asymmetric, irrational-looking
    nodes["nvl_bw"] = np.where(frac[:, None, None], nodes["nvl_bw"] * jitter, nodes["nvl_bw"])
    nodes["nvl_bw"][5] *= 0.0                       # active links of zero bandwidth never extend a group
    pods = synth.make_pods(257, seed=21, k_lo=2, pref_probs=(0, 0.8, 0.2, 0, 0), minbw_choices=(0.0, 500.0))
    led = synth.make_ledger(nodes, 0.3, seed=21)
    eng = engine_for(kg, nodes, led)
    s, top = eng.score_batch(pods, len(nodes), want_scores=True, topk=4)
    o, o_top, _ = orc.score_batch(abi, nodes, led, pods, want_scores=True, topk=4, threads=4)
    assert s.tobytes() == o.tobytes()
    assert_same_placements(top, o_top)
    dec = eng.schedule_queue(pods)
    o_dec, _, _ = orc.schedule_queue(abi, nodes, led, pods)
    assert_same_placements(dec, o_dec)
    eng.close()


def test_gang_scheduling_extension(kg, orc):
    """E5: config 5 shape (8 members x 8 GPUs per gang) on a cluster too small for all gangs, so
    some gangs roll back in the middle of the queue; plus hand-built mixed gangs."""
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(5, 0.012)            # 120 nodes, 15 gangs of 8 x 8 GPUs
    gang_of = (np.arange(len(pods)) // 8).astype(np.uint32)
    eng = engine_for(kg, nodes, led)
    dec = eng.schedule_gangs(pods, gang_of)
    o_dec, o_led = orc.schedule_gangs(abi, nodes, led, pods, gang_of)
    assert_same_placements(dec, o_dec)
    assert eng.stats()["n_allocations"] == len(o_led)
    assert (dec["node_idx"] == abi.NO_NODE).any() and (dec["node_idx"] != abi.NO_NODE).any()
    # the image must match the oracle's final ledger (rolled-back commits really gone)
    s, _ = eng.score_batch(pods[:16], len(nodes), want_scores=True)
    o, _, _ = orc.score_batch(abi, nodes, o_led, pods[:16], want_scores=True)
    assert s.tobytes() == o.tobytes()
    eng.close()
    # ragged gangs with mixed sizes and preferences, several failures in a row
    nodes = synth.make_cluster(40, "mixed", seed=3)
    pods = synth.make_pods(300, seed=3, k_lo=3, k_hi=8)
    sizes = synth.SplitMix64(5).randint(1, 9, 100)
    gang_of = np.repeat(np.arange(100), sizes)[:300].astype(np.uint32)
    eng = engine_for(kg, nodes)
    dec = eng.schedule_gangs(pods, gang_of)
    o_dec, o_led = orc.schedule_gangs(abi, nodes, None, pods, gang_of)
    assert_same_placements(dec, o_dec)
    assert (dec["flags"] & abi.F_GANG_ROLLBACK).any()
    eng.close()


def test_side_terms_extension(kg, orc):
    """E4: bisection bandwidth / ring bottleneck of placements, GPU vs brute force (the only
    floating-point accumulation is the cut sum, in a fixed order: bit-exact expected)."""
    abi, synth = kg.abi, kg.synth
    nodes = synth.make_cluster(300, "mixed", seed=8, inactive_link_frac=0.08)
    rng = synth.SplitMix64(9)
    jitter = 0.5 + rng.uniform(len(nodes) * 64).reshape(len(nodes), 8, 8)
    nodes["nvl_bw"] = nodes["nvl_bw"] * jitter                     # asymmetric, non-integral
    pods = synth.make_pods(400, seed=8)
    eng = engine_for(kg, nodes)
    dec = eng.schedule_queue(pods)
    extra = np.zeros(3, dtype=abi.PLACEMENT)
    extra["node_idx"] = [abi.NO_NODE, 0, 5]
    extra["gpu_mask"] = [0, 0xFF, 0x00]
    pls = np.concatenate([dec, extra])
    t = eng.placement_terms(pls)
    o = orc.placement_terms(abi, nodes, pls)
    np.testing.assert_allclose(t["bisection_gbps"], o["bisection_gbps"], rtol=REL_TOL, atol=0)
    np.testing.assert_allclose(t["ring_bottleneck_gbps"], o["ring_bottleneck_gbps"], rtol=REL_TOL, atol=0)
    assert t.tobytes() == o.tobytes()
    assert (t["bisection_gbps"] > 0).any() and (t["ring_bottleneck_gbps"] > 0).any()
    eng.close()


def test_mig_pack_and_rebalance_extension(kg, orc):
    """E2: GPU bin-packing kernel and rebalancing pass vs the plain-loop oracle; config 4 at
    full size (10,000 GPUs, 50,000 slice requests) through size-independent properties."""
    abi, synth = kg.abi, kg.synth
    eng = kg.Engine(max_nodes=1)
    gpus, profiles = synth.make_mig(1500, 9000)
    a, g_after = eng.mig_pack(gpus, profiles)
    o, og_after = orc.mig_pack(abi, gpus, profiles)
    assert a.tobytes() == o.tobytes() and g_after.tobytes() == og_after.tobytes()
    assert (a["gpu_idx"] == abi.NO_NODE).any() and (a["gpu_idx"] != abi.NO_NODE).sum() > 4000
    target = np.zeros(10)
    target[[0, 2, 4]] = [0.5, 0.3, 0.2]
    target[[1, 3, 5, 9]] = [0.7, 0.2, 0.07, 0.03]
    target /= 2.0
    for tbl in (gpus, g_after):
        ops, supply, n = eng.mig_rebalance(tbl, target)
        o_ops, o_supply, o_n = orc.mig_rebalance(abi, tbl, target)
        assert n == o_n and ops.tobytes() == o_ops.tobytes() and supply.tolist() == o_supply.tolist()
    ops, supply, n = eng.mig_rebalance(gpus, target, ops_cap=10)       # truncated op buffer still reports the full plan size
    assert len(ops) == 10 and n == o_n - 0 or n > 10
    with pytest.raises(kg.EngineError):
        eng.mig_rebalance(gpus, np.full(10, 0.2))                     # sums to 2.0 > 1.01 (mig_controller.go:263-270)
    # edge cases: no GPUs, no requests, invalid profile ids
    a, _ = eng.mig_pack(gpus[:0], profiles[:5])
    assert (a["gpu_idx"] == abi.NO_NODE).all()
    a, _ = eng.mig_pack(gpus, profiles[:0])
    assert len(a) == 0
    a, _ = eng.mig_pack(gpus[:10], np.array([0, 11, 200], dtype=np.uint8))
    assert (a["gpu_idx"] == abi.NO_NODE).all()
    # full size: every assignment is a legal, non-overlapping placement on an enabled GPU of the right class
    gpus, profiles = synth.make_mig(10000, 50000)
    a, g_after = eng.mig_pack(gpus, profiles)
    occ = gpus["occ"].astype(np.int64).copy()
    ok = a["gpu_idx"] != abi.NO_NODE
    legal_start = {1: range(0, 7), 2: (0, 2, 4), 4: (0, 4), 8: (0,)}
    for r in np.nonzero(ok)[0][::7]:
        x = a[r]
        assert int(x["start"]) in legal_start[int(x["size"])]
    for x in a[ok]:
        m = ((1 << int(x["size"])) - 1) << int(x["start"])
        gidx = int(x["gpu_idx"])
        assert occ[gidx] & m == 0 and gpus["enabled"][gidx] and gpus["cls"][gidx] == (int(x["profile"]) - 1) % 2
        occ[gidx] |= m
    assert np.array_equal(occ.astype(np.uint8), g_after["occ"])
    # a failed request really had no fit at that moment is covered by the oracle parity above
    eng.close()


def test_mig_large_table_and_reconfigure(kg, orc):
    """E2: a table too large for the tournament-tree kernel takes the scan form of the packer (vs the oracle); the
    reconfiguration pass with destroy ops vs its plain-loop oracle on random instance layouts, config 4's size included."""
    abi, synth = kg.abi, kg.synth
    eng = kg.Engine(max_nodes=1)
    gpus, profiles = synth.make_mig(26000, 1200, seed=31)
    a, after = eng.mig_pack(gpus, profiles)
    o, o_after = orc.mig_pack(abi, gpus, profiles)
    assert a.tobytes() == o.tobytes() and after.tobytes() == o_after.tobytes()
    rng = synth.SplitMix64(5)
    for n in (1, 37, 1024, 1025, 10000):
        st = np.zeros(n, dtype=abi.MIG_GPU_STATE)
        st["cls"] = rng.randint(0, 1, n); st["enabled"] = (rng.uniform(n) > 0.05).astype(np.uint8)
        st["created"] = (rng.randint(0, (1 << 14) - 1, n) & rng.randint(0, (1 << 14) - 1, n)).astype(np.uint16)
        busy = rng.uniform(n) < 0.4
        st["in_use"] = np.where(busy, st["created"] & rng.randint(0, (1 << 14) - 1, n).astype(np.uint16), 0).astype(np.uint16)
        t = rng.uniform(10) * (rng.uniform(10) < 0.5)
        target = t / t.sum() * 0.9 if t.sum() > 0 else t
        ops, supply, n_ops = eng.mig_reconfigure(st, target)
        o_ops, o_supply, o_n = orc.mig_reconfigure(abi, st, target)
        assert n_ops == o_n and supply.tolist() == o_supply.tolist(), n
        assert ops.tobytes() == o_ops.tobytes(), n
        ops2, _, n2 = eng.mig_reconfigure(st, target, ops_cap=5)            # a short op buffer still reports the full plan
        assert n2 == o_n and ops2.tobytes() == o_ops[:5].tobytes()
    bad = np.zeros(1, dtype=abi.MIG_GPU_STATE); bad["in_use"] = 1
    with pytest.raises(kg.EngineError):
        eng.mig_reconfigure(bad, np.zeros(10))
    eng.close()


def test_latency_path_small_batches(kg, orc):
    """Per-pod Prioritize calls (<= 32 pods, top-1 only) take the zero-copy single-launch path:
    pods of different preferences side by side (lane groups of one warp on different branches),
    with and without hints, after incremental updates, many calls in a row."""
    abi, synth = kg.abi, kg.synth
    for num, scale in ((1, 1.0), (2, 0.15), (3, 0.03)):
        nodes, pods, led = synth.config(num, scale)
        pods = pods[:96].copy()
        pods["topology_pref"] = np.arange(len(pods)) % 5                   # every preference, interleaved
        pods["gpu_count"] = 1 + (np.arange(len(pods)) * 7) % 9            # 1..9 (9 never fits)
        hint = synth.make_hint(len(pods), len(nodes), 0.05)
        eng = engine_for(kg, nodes, led)
        _, want, _ = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=False, topk=1)
        _, want_nohint, _ = orc.score_batch(abi, nodes, led, pods, want_scores=False, topk=1)
        for size in (1, 3, 8, 32):
            for p0 in range(0, len(pods) - size + 1, size):
                _, got = eng.score_batch(pods[p0:p0 + size], len(nodes), hint=hint[p0:p0 + size], want_scores=False, topk=1)
                assert_same_placements(got, want[p0:p0 + size])
        _, got = eng.score_batch(pods[:17], len(nodes), want_scores=False, topk=1)
        assert_same_placements(got, want_nohint[:17])
        # a commit between two calls is seen by the next call
        eng.commit(999_999, 1, 0x0F)
        led2 = np.concatenate([led, ledger_of(abi, [(999_999, 1, 0x0F)])])
        _, got = eng.score_batch(pods[:9], len(nodes), want_scores=False, topk=1)
        _, want2, _ = orc.score_batch(abi, nodes, led2, pods[:9], want_scores=False, topk=1)
        assert_same_placements(got, want2)
        eng.close()


def test_topology_hint_and_gauges(kg, orc):
    """The reference's second node-parallel scorer (DiscoveryService.GetTopologyHint) and the
    exporter's per-node gauge, GPU vs the restatement, incl. the hand-derived cases."""
    abi, synth = kg.abi, kg.synth
    nodes = synth.make_cluster(3000, "mixed", seed=31, unhealthy_frac=0.1, inactive_link_frac=0.05, partial_node_frac=0.2, mig_frac=0.4)
    reqs = synth.make_hint_reqs(500, seed=31)
    eng = engine_for(kg, nodes)
    got = eng.topology_hints(reqs)
    want = orc.topology_hints(abi, nodes, reqs)
    for f in abi.TOPOLOGY_HINT.names:
        if f == "bandwidth_estimate":
            np.testing.assert_allclose(got[f], want[f], rtol=REL_TOL, atol=0)
        else:
            assert np.array_equal(got[f], want[f]), f
    assert got.tobytes() == want.tobytes()
    assert (got["node_idx"] == abi.NO_NODE).any() and (got["score"] >= 90).any()
    assert np.array_equal(eng.topology_gauges(len(nodes)), orc.topology_gauges(abi, nodes))
    # an informer update changes the answer
    nodes2 = nodes.copy()
    best = int(got[got["node_idx"] != abi.NO_NODE][0]["node_idx"])
    nodes2[best]["healthy_mask"] = 0
    eng.upsert_node(best, nodes2[best])
    assert eng.topology_hints(reqs).tobytes() == orc.topology_hints(abi, nodes2, reqs).tobytes()
    # sharded engine: global indices
    e2 = kg.Engine(max_nodes=1000, node_base=2000)
    e2.upsert_nodes(nodes[2000:])
    cfg = abi.default_config(1000, node_base=2000)
    assert e2.topology_hints(reqs[:50]).tobytes() == orc.topology_hints(abi, nodes[2000:], reqs[:50], cfg=cfg).tobytes()
    assert eng.topology_hints(reqs[:0]).shape == (0,)
    eng.close(); e2.close()


def test_non_default_config(kg, orc):
    """SchedulerConfig weights, hint bonus and the NVLink normalisation are engine parameters
    (types.go:346-392, scheduler.go:282-286, 368): non-default values, incl. zero and negative weights."""
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(2, 0.1)
    pods = synth.make_pods(120, seed=77, arch_frac=0.2)
    hint = synth.make_hint(len(pods), len(nodes), 0.1)
    for wt, wr, wb, bonus, norm in ((60, 30, 10, 25, 600.0), (0, 100, 0, 0, 900.0), (100, -20, 50, 7, 1800.0), (33, 33, 34, 1000, 450.5)):
        cfg = abi.default_config(len(nodes))
        cfg.topology_weight, cfg.resource_weight, cfg.balance_weight, cfg.hint_bonus, cfg.nvlink_norm_gbps = wt, wr, wb, bonus, norm
        eng = kg.Engine(max_nodes=len(nodes), config=cfg)
        eng.upsert_nodes(nodes)
        eng.seed_ledger(led)
        s, t = eng.score_batch(pods, len(nodes), hint=hint, want_scores=True, topk=3)
        o, ot, _ = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=True, topk=3, cfg=cfg)
        assert s.tobytes() == o.tobytes()
        assert_same_placements(t, ot)
        d = eng.schedule_queue(pods, hint=hint)
        od, _, _ = orc.schedule_queue(abi, nodes, led, pods, hint=hint, cfg=cfg)
        assert_same_placements(d, od)
        eng.close()


def test_concurrent_callers_share_one_engine(kg, orc):
    """Entry points take the engine's mutex (mirrors s.mu, scheduler.go:17): threads hammering
    stateless scoring while another thread commits and releases must all get consistent answers."""
    import threading
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(2, 0.05)
    pods = pods[:40]
    eng = engine_for(kg, nodes, led)
    _, base, _ = orc.score_batch(abi, nodes, led, pods, want_scores=False, topk=1)
    spare = int(np.setdiff1d(np.arange(len(nodes)), led["node_idx"])[0])        # a node the ledger does not touch (a commit on taken GPUs is refused)
    led2 = np.concatenate([led, ledger_of(abi, [(77_000_001, spare, 0xFF)])])
    _, alt, _ = orc.score_batch(abi, nodes, led2, pods, want_scores=False, topk=1)
    errors = []

    def reader():
        try:
            for _ in range(60):
                _, t = eng.score_batch(pods, len(nodes), want_scores=False, topk=1)
                if t.tobytes() not in (base.tobytes(), alt.tobytes()):
                    errors.append("torn read")
        except Exception as ex:           # noqa: BLE001
            errors.append(repr(ex))

    def writer():
        try:
            for _ in range(60):
                eng.commit(77_000_001, spare, 0xFF)
                eng.release(77_000_001)
        except Exception as ex:           # noqa: BLE001
            errors.append(repr(ex))

    threads = [threading.Thread(target=reader) for _ in range(3)] + [threading.Thread(target=writer)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors[:3]
    _, t = eng.score_batch(pods, len(nodes), want_scores=False, topk=1)
    assert t.tobytes() == base.tobytes()
    eng.close()


def test_topk_merge_device_equals_host_merge(kg):
    """The merge every rank runs after the exchange: device kernel vs kgwe_topk_merge_host on random ranked lists
    (ties on score across shards, short lists, empty shards)."""
    import torch
    abi, synth = kg.abi, kg.synth
    rng = synth.SplitMix64(404)
    R, P, K = 5, 300, 6
    lists = np.zeros((R, P, K), dtype=abi.PLACEMENT)
    lists["node_idx"] = abi.NO_NODE
    for r in range(R):
        n_valid = rng.randint(0, K, P)
        totals = np.sort(rng.randint(1, 12, P * K).reshape(P, K), axis=1)[:, ::-1]          # few distinct scores -> many ties
        nodes_ = np.sort(rng.randint(0, 199, P * K).reshape(P, K), axis=1) + 1000 * r       # ascending inside a shard, disjoint across shards
        # canonical order inside a list: score desc, node asc -> sort node ascending within equal scores
        for p in range(P):
            order = np.lexsort((nodes_[p], -totals[p]))
            t, n = totals[p][order], nodes_[p][order]
            v = int(n_valid[p])
            lists[r, p, :v]["total"] = t[:v]
            lists[r, p, :v]["node_idx"] = n[:v]
            lists[r, p, :v]["gpu_mask"] = (n[:v] % 255).astype(np.uint8)
    want = kg.sharded.merge_host(lists, K)
    eng = kg.Engine(max_nodes=1)
    d_lists = torch.from_numpy(lists.view(np.uint8).reshape(R, P, K, abi.PLACEMENT.itemsize)).cuda()
    d_out = torch.zeros((P, K, abi.PLACEMENT.itemsize), dtype=torch.uint8, device="cuda")
    eng.topk_merge_device(d_lists.data_ptr(), R, P, K, d_out.data_ptr())
    torch.cuda.synchronize()
    got = d_out.cpu().numpy().view(abi.PLACEMENT).reshape(P, K)
    assert got.tobytes() == want.tobytes()
    # reference semantics of the merge itself: global (score desc, node asc) order of the union
    for p in (0, 17, 299):
        union = [(int(x["total"]), int(x["node_idx"])) for r in range(R) for x in lists[r, p] if x["node_idx"] != abi.NO_NODE]
        union.sort(key=lambda x: (-x[0], x[1]))
        exp = union[:K]
        assert [(int(x["total"]), int(x["node_idx"])) for x in want[p] if x["node_idx"] != abi.NO_NODE] == exp
    eng.close()


def test_snapshot_and_restore(kg, orc):
    """Checkpoint / resume: image + ledger saved after a queue has run, loaded into a fresh engine, and both
    engines must keep answering identically (stateless scores, further scheduling, releases)."""
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(2, 0.08)
    eng = engine_for(kg, nodes, led)
    eng.schedule_queue(pods[:40])
    eng.remove_node(5)
    blob = eng.snapshot()
    assert blob == eng.snapshot()                                   # deterministic
    fresh = kg.Engine(max_nodes=len(nodes) + 10)
    fresh.restore(blob)
    assert fresh.stats()["n_allocations"] == eng.stats()["n_allocations"] and fresh.stats()["n_nodes"] == eng.stats()["n_nodes"]
    a, ta = eng.score_batch(pods[40:70], len(nodes), want_scores=True, topk=3)
    b, tb = fresh.score_batch(pods[40:70], len(nodes), want_scores=True, topk=3)
    assert a.tobytes() == b.tobytes() and ta.tobytes() == tb.tobytes()
    da, db = eng.schedule_queue(pods[40:70]), fresh.schedule_queue(pods[40:70])
    assert da.tobytes() == db.tobytes()
    wl = int(pods[41]["wl_id"])
    if da[1]["node_idx"] != abi.NO_NODE:
        eng.release(wl); fresh.release(wl)
    a, _ = eng.score_batch(pods[:10], len(nodes), want_scores=True)
    b, _ = fresh.score_batch(pods[:10], len(nodes), want_scores=True)
    assert a.tobytes() == b.tobytes()
    small = kg.Engine(max_nodes=4)
    with pytest.raises(kg.EngineError):
        small.restore(blob)                                         # too many nodes for this engine
    with pytest.raises(kg.EngineError):
        fresh.restore(blob[:100])
    # a corrupted blob (a record claiming 200 GPUs; a ledger entry naming a node far away) is rejected whole,
    # and the engine keeps the state it had
    before = fresh.snapshot()
    bad = bytearray(blob)
    n_gpus_off = 32 + 2 * 768 + 64 * 8 + 3 * 64 + 5 * 8 + abi.MIG_PROFILES       # header, record 2, up to n_gpus
    bad[n_gpus_off] = 200
    with pytest.raises(kg.EngineError):
        fresh.restore(bytes(bad))
    bad = bytearray(blob)
    bad[-8:-4] = (0x7FFFFFFF).to_bytes(4, "little")                                # last ledger entry's node
    with pytest.raises(kg.EngineError):
        fresh.restore(bytes(bad))
    assert fresh.snapshot() == before
    eng.close(); fresh.close(); small.close()


def test_differential_fuzz_few_seeds(kg, orc):
    """tests/fuzz_parity.py for a handful of seeds: every feature at random (ragged nodes, inactive / zero /
    fractional / irregular links, MIG, hints, architectures, min-bandwidth, all preferences, counts 0..9,
    exhaustive flags, both kernel-variant flags), engine vs oracle, stateless + latency path + queue."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("fuzz_parity", os.path.join(os.path.dirname(os.path.abspath(__file__)), "fuzz_parity.py"))
    fz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fz)
    for seed in (101, 102, 103, 104, 105, 106):
        fz.one_seed(kg, orc, seed)
    # and the stateful life cycle: informer events, ledger calls, queues and scoring interleaved at random on one
    # engine, mirrored on a plain model that the oracle is given after every step; checkpoint at the end
    for seed in (201, 202, 203, 204):
        fz.life_cycle_seed(kg, orc, seed)
    # and E2 on random MIG tables (any occupancy byte, every profile id, random targets)
    for seed in (301, 302, 303, 304, 305, 306):
        fz.mig_seed(kg, orc, seed)
