"""Three restatements of the reference's scoring path must agree: the two builds of the C++ oracle (faithful, indexed) and
the pure-Python one in oracle/restatement.py, which was written from scheduler.go on its own with the reference's data
shapes (UUID strings, adjacency dicts, a dict ledger).  Small random clusters with every reference feature: ragged nodes,
unhealthy GPUs, inactive / missing / zero / fractional / asymmetric links, NUMA and PCIe groups, architectures, MIG counts,
a pre-seeded ledger, optimizer hints, all preferences, GPUCount 0..9; stateless matrix and the stateful queue."""
import numpy as np
import pytest

from oracle import restatement as rs


def _case(kg, seed):
    synth, abi = kg.synth, kg.abi
    rng = synth.SplitMix64(seed * 31 + 5)
    u = rng.uniform(8)
    n_nodes, n_pods = int(10 + u[0] * 25), int(20 + u[1] * 30)
    nodes = synth.make_cluster(n_nodes, "mixed" if u[2] < 0.7 else "dgx", seed=seed, unhealthy_frac=0.15, inactive_link_frac=0.15,
                               partial_node_frac=0.4, mig_frac=0.4)
    kind = seed % 4
    if kind == 1:        # fractional, asymmetric
        nodes["nvl_bw"] = nodes["nvl_bw"] * (0.5 + rng.uniform(n_nodes * 64).reshape(n_nodes, 8, 8))
    elif kind == 2:      # whole-number, irregular
        nodes["nvl_bw"] = np.where(nodes["nvl_bw"] > 0, rng.randint(0, 1500, n_nodes * 64).reshape(n_nodes, 8, 8).astype(np.float64), 0.0)
    elif kind == 3:      # some links reported with zero bandwidth, some directions missing
        drop = rng.uniform(n_nodes * 64).reshape(n_nodes, 8, 8) < 0.1
        nodes["nvl_bw"] = np.where(drop, 0.0, nodes["nvl_bw"])
        for i in range(n_nodes):
            g = int(rng.randint(0, 7, 1)[0])
            nodes["nvl_present"][i][g] &= ~np.uint8(1 << int(rng.randint(0, 7, 1)[0]))
            nodes["nvl_active"][i] &= nodes["nvl_present"][i]
    pods = synth.make_pods(n_pods, seed=seed, k_lo=0, k_hi=9, pref_probs=(0.15, 0.3, 0.15, 0.2, 0.2), arch_frac=0.3, mig_frac=0.3)
    pods["flags"] = 0
    pods["min_bandwidth_gbps"] = 0.0
    ledger = synth.make_ledger(nodes, 0.3, seed=seed)
    hint = synth.make_hint(n_pods, n_nodes, 0.1, seed=seed)
    return nodes, pods, ledger, hint


def _order_list(order):
    n = (int(order) >> 24) & 15
    return [(int(order) >> (3 * i)) & 7 for i in range(n)]


@pytest.mark.parametrize("seed", range(1, 41))
def test_score_matrix(kg, orc, seed):
    abi = kg.abi
    nodes, pods, ledger, hint = _case(kg, seed)
    N, P = len(nodes), len(pods)
    faithful, f_top, _ = orc.score_batch(abi, nodes, ledger, pods, hint, want_scores=True, topk=1, mode=orc.FAITHFUL)
    indexed, i_top, _ = orc.score_batch(abi, nodes, ledger, pods, hint, want_scores=True, topk=1, mode=orc.INDEXED)
    assert faithful.tobytes() == indexed.tobytes() and f_top.tobytes() == i_top.tobytes()
    sched = rs.Scheduler(rs.nodes_from_images(nodes))
    rs.seed_ledger(sched, ledger)
    bits = np.unpackbits(hint, axis=1, bitorder="little")[:, :N]
    for p in range(P):
        w = rs.workload_from_pod(pods[p])
        preferred = {f"node{n}" for n in np.flatnonzero(bits[p])}
        scores = sched.score_nodes(w, preferred)
        best = None
        for n, s in enumerate(scores):
            rec = indexed[p, n]
            if s is None:
                assert rec.tobytes() == bytes(16), f"seed {seed} pod {p} node {n}: the Python restatement filters the node out"
                continue
            assert int(rec["flags"]) & abi.F_ELIGIBLE, f"seed {seed} pod {p} node {n}: eligible only in the Python restatement"
            got = (int(rec["total"]), int(rec["topology"]), int(rec["resource"]), int(rec["balance"]), _order_list(rec["order"]),
                   bool(int(rec["flags"]) & abi.F_KEPT), bool(int(rec["flags"]) & abi.F_HINT))
            want = (s.total, s.topology, s.resource, s.balance, rs.gpu_indices(s.selected), s.total > 0, s.hinted)
            assert got == want, f"seed {seed} pod {p} node {n}: {got} != {want}"
            assert int(rec["gpu_mask"]) == sum(1 << g for g in rs.gpu_indices(s.selected))
            if s.total > 0 and (best is None or s.total > best.total):
                best = s
        t = i_top[p, 0]
        if best is None:
            assert int(t["node_idx"]) == abi.NO_NODE
        else:
            assert int(t["node_idx"]) == best.node.index and int(t["total"]) == best.total
            assert abs(float(t["est_bandwidth_gbps"]) - sched.estimate_bandwidth(best)) <= 1e-9 * max(1.0, abs(float(t["est_bandwidth_gbps"])))


@pytest.mark.parametrize("seed", range(101, 121))
def test_stateful_queue(kg, orc, seed):
    abi = kg.abi
    nodes, pods, ledger, hint = _case(kg, seed)
    N = len(nodes)
    pods["wl_id"][len(pods) // 2:len(pods) // 2 + 4] = pods["wl_id"][:4]          # scheduled again: the same UID overwrites (:699)
    want, led_after, _ = orc.schedule_queue(abi, nodes, ledger, pods, hint, mode=orc.FAITHFUL)
    sched = rs.Scheduler(rs.nodes_from_images(nodes))
    rs.seed_ledger(sched, ledger)
    bits = np.unpackbits(hint, axis=1, bitorder="little")[:, :N]
    for p in range(len(pods)):
        best, bw = sched.schedule(rs.workload_from_pod(pods[p]), {f"node{n}" for n in np.flatnonzero(bits[p])})
        d = want[p]
        if best is None:
            assert int(d["node_idx"]) == abi.NO_NODE, f"seed {seed} pod {p}"
            continue
        assert (int(d["node_idx"]), int(d["total"]), int(d["topology"]), int(d["resource"]), int(d["balance"]), _order_list(d["order"])) == \
               (best.node.index, best.total, best.topology, best.resource, best.balance, rs.gpu_indices(best.selected)), f"seed {seed} pod {p}"
        assert abs(float(d["est_bandwidth_gbps"]) - bw) <= 1e-9 * max(1.0, abs(bw))
    assert len(sched.allocations) == len(led_after)


@pytest.mark.parametrize("seed", range(201, 221))
def test_topology_hint_and_gauge(kg, orc, seed):
    """DiscoveryService.GetTopologyHint (discovery.go:222-247, 378-541) and the exporter's per-node gauge
    (prometheus_exporter.go:517-539): the C++ oracle against the Python restatement."""
    abi = kg.abi
    nodes, _, _, _ = _case(kg, seed)
    reqs = kg.synth.make_hint_reqs(40, seed=seed)
    want = orc.topology_hints(abi, nodes, reqs)
    objs = rs.nodes_from_images(nodes)
    disc = rs.Discovery(objs)
    for i, rq in enumerate(reqs):
        h = disc.get_topology_hint(rs.HintRequirements(int(rq["count"]), int(rq["min_memory_gb"]), bool(rq["prefer_nvlink"]), bool(rq["prefer_same_numa"]), int(rq["arch"])))
        w = want[i]
        if h is None:
            assert int(w["node_idx"]) == abi.NO_NODE and int(w["score"]) == 0, f"seed {seed} request {i}"
            continue
        n = (int(w["order"]) >> 24) & 15
        order = [(int(w["order"]) >> (3 * j)) & 7 for j in range(n)]
        assert (int(w["node_idx"]), int(w["score"]), order) == (h[0], h[1], rs.gpu_indices(h[2])), f"seed {seed} request {i}"
        assert abs(float(w["bandwidth_estimate"]) - h[3]) <= 1e-9 * max(1.0, abs(h[3]))
    gauges = orc.topology_gauges(abi, nodes)
    assert [int(x) for x in gauges] == [rs.exporter_topology_score(n) for n in objs]
