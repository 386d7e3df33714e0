"""Differential fuzzing: random clusters / pod queues with every feature switched on at random, engine vs
oracle (indexed build), stateless matrix + top-k and the stateful queue, over many seeds.
  python tests/fuzz_parity.py [n_seeds] [first_seed]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry


def make_case(abi, synth, seed, n_lo=40, n_span=400, p_lo=40, p_span=200):
    """One random case: (nodes, pods, ledger, hint or None, engine flags, top-k, gang_of, bandwidth kind)."""
    rng = synth.SplitMix64(seed * 7919 + 13)
    u = rng.uniform(16)
    N = int(n_lo + u[0] * n_span)
    P = int(p_lo + u[1] * p_span)
    nodes = synth.make_cluster(N, "mixed" if u[2] < 0.8 else "dgx", seed=seed, unhealthy_frac=u[3] * 0.3, inactive_link_frac=u[4] * 0.2,
                               partial_node_frac=u[5] * 0.5, mig_frac=u[6])
    kind = int(u[7] * 6)
    if kind == 1:      # fractional, asymmetric bandwidths on some nodes (FP64 path)
        frac = rng.uniform(N) < 0.4
        jit = 0.5 + rng.uniform(N * 64).reshape(N, 8, 8)
        nodes["nvl_bw"] = np.where(frac[:, None, None], nodes["nvl_bw"] * jit, nodes["nvl_bw"])
    elif kind == 2:    # whole-number but irregular, asymmetric bandwidths (integer path, no common unit)
        jit = rng.randint(1, 2000, N * 64).reshape(N, 8, 8).astype(np.float64)
        nodes["nvl_bw"] = np.where(nodes["nvl_bw"] > 0, jit, 0.0)
    elif kind == 5:    # one fractional bandwidth per node on most nodes (uniform walk with float64 totals), per-link jitter on the rest
        per_node = 0.5 + rng.uniform(N)
        jit = 0.5 + rng.uniform(N * 64).reshape(N, 8, 8)
        odd = rng.uniform(N) < 0.25
        nodes["nvl_bw"] = nodes["nvl_bw"] * np.where(odd[:, None, None], jit, per_node[:, None, None])
    elif kind == 3:    # zero-bandwidth links here and there
        z = rng.uniform(N * 64).reshape(N, 8, 8) < 0.15
        nodes["nvl_bw"] = np.where(z, 0.0, nodes["nvl_bw"])
    nodes["numa"] = rng.randint(-2, 2, N * 8).reshape(N, 8).astype(np.int8)
    nodes["pcie_switch"] = rng.randint(0, 3, N * 8).reshape(N, 8).astype(np.uint8)
    nodes["util"] = np.floor(rng.uniform(N * 8).reshape(N, 8) * 400) / 4.0
    if kind == 4:      # values at the edges of what an upsert accepts
        bw_set = np.array([0.0, 1.0, 0.5, 100000.0, 99999.75, 3.0000000001, 1e-300, 899.9999999999999, 900.0])
        nodes["nvl_bw"] = np.where(nodes["nvl_bw"] > 0, bw_set[rng.randint(0, len(bw_set) - 1, N * 64)].reshape(N, 8, 8), 0.0)
        pc_set = np.array([0.0, 100000.0, 15.75, 64.0])
        nodes["pcie_bw"] = pc_set[rng.randint(0, len(pc_set) - 1, N * 8)].reshape(N, 8)
        ut_set = np.array([-50.0, 0.0, 29.999999999999996, 30.0, 59.99999999999999, 60.0, 1e300, 100.0])
        nodes["util"] = ut_set[rng.randint(0, len(ut_set) - 1, N * 8)].reshape(N, 8)
        fb_set = np.array([0, 2**64 - 1, 2**63, 16 * 2**30, 16 * 2**30 - 1, 2**61], dtype=np.uint64)
        nodes["free_bytes"] = fb_set[rng.randint(0, len(fb_set) - 1, N * 8)].reshape(N, 8)
        nodes["numa"] = np.array([-128, 127, 0, -1], dtype=np.int8)[rng.randint(0, 3, N * 8)].reshape(N, 8)
        nodes["pcie_switch"] = np.array([0, 255, 7], dtype=np.uint8)[rng.randint(0, 2, N * 8)].reshape(N, 8)
        nodes["mig_free"] = np.array([0, 1, 255], dtype=np.uint8)[rng.randint(0, 2, N * abi.MIG_PROFILES)].reshape(N, abi.MIG_PROFILES)
    prefs = rng.uniform(5) + 0.05
    pods = synth.make_pods(P, seed=seed, k_lo=0, k_hi=9, pref_probs=tuple(prefs / prefs.sum()), arch_frac=u[8] * 0.6,
                           minbw_choices=(0.0, 0.0, 200.0, 650.0, 1000.0), mig_frac=u[9] * 0.5)
    pods["min_memory_gb"] = rng.randint(0, 200, P)
    if kind == 4:
        pods["min_memory_gb"] = np.array([0, 1, 16, 2**20, 2**19], dtype=np.int64)[rng.randint(0, 4, P)]
        pods["min_bandwidth_gbps"] = np.array([0.0, 1e-300, 64.0, 99999.75, 1e300])[rng.randint(0, 4, P)]
        pods["mig_count"] = np.where(pods["mig_profile"] != abi.MIG_NONE, np.array([0, 1, 255, 256, -1])[rng.randint(0, 4, P)], 0)
    exh = rng.uniform(P) < u[10] * 0.5
    pods["flags"] = np.where(exh, abi.POD_EXHAUSTIVE, 0).astype(np.uint8)
    led = synth.make_ledger(nodes, u[11] * 0.8, seed=seed)
    hint = synth.make_hint(P, N, u[12] * 0.1, seed=seed) if u[13] < 0.5 else None
    flags = [0, abi.CFG_MEMOISE_GROUPS, abi.CFG_SUBSET_TABLES, abi.CFG_MEMOISE_GROUPS | abi.CFG_SUBSET_TABLES][int(u[14] * 4)]
    topk = [1, 2, 5, 16][int(u[15] * 4)]
    sizes = rng.randint(1, 9, P)
    gang_of = np.repeat(np.arange(P), sizes)[:P].astype(np.uint32)
    return nodes, pods, led, hint, flags, topk, gang_of, kind


def one_seed(kg, orc, seed):
    abi, synth = kg.abi, kg.synth
    nodes, pods, led, hint, flags, topk, gang_of, kind = make_case(abi, synth, seed)
    N, P = len(nodes), len(pods)
    eng = kg.Engine(max_nodes=N, flags=flags)
    eng.upsert_nodes(nodes)
    eng.seed_ledger(led)
    s, t = eng.score_batch(pods, N, hint=hint, want_scores=True, topk=topk)
    o, ot, _ = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=True, topk=topk, threads=4)
    assert s.tobytes() == o.tobytes(), f"seed {seed}: score matrix differs"
    assert t.tobytes() == ot.tobytes(), f"seed {seed}: top-{topk} differs"
    for p0 in range(0, min(P, 24), 3):       # latency path
        _, t1 = eng.score_batch(pods[p0:p0 + 3], N, hint=None if hint is None else hint[p0:p0 + 3], want_scores=False, topk=1)
        _, o1, _ = orc.score_batch(abi, nodes, led, pods[p0:p0 + 3], hint=None if hint is None else hint[p0:p0 + 3], want_scores=False, topk=1)
        assert t1.tobytes() == o1.tobytes(), f"seed {seed}: latency path differs at pod {p0}"
    d = eng.schedule_queue(pods, hint=hint)
    od, oled, _ = orc.schedule_queue(abi, nodes, led, pods, hint=hint)
    assert d.tobytes() == od.tobytes(), f"seed {seed}: queue decisions differ"
    s2, _ = eng.score_batch(pods[:8], N, want_scores=True)
    o2, _, _ = orc.score_batch(abi, nodes, oled, pods[:8], want_scores=True)
    assert s2.tobytes() == o2.tobytes(), f"seed {seed}: image after the queue differs"
    t = eng.placement_terms(d)
    ot = orc.placement_terms(abi, nodes, d)
    assert t.tobytes() == ot.tobytes(), f"seed {seed}: side terms differ"
    reqs = synth.make_hint_reqs(64, seed=seed)
    assert eng.topology_hints(reqs).tobytes() == orc.topology_hints(abi, nodes, reqs).tobytes(), f"seed {seed}: topology hints differ"
    assert np.array_equal(eng.topology_gauges(N), orc.topology_gauges(abi, nodes)), f"seed {seed}: gauges differ"
    eng.close()
    # gangs on a fresh engine: ragged gang sizes, all-or-nothing with rollbacks
    eng = kg.Engine(max_nodes=N, flags=flags)
    eng.upsert_nodes(nodes)
    eng.seed_ledger(led)
    g = eng.schedule_gangs(pods, gang_of, hint=hint)
    og, ogl = orc.schedule_gangs(abi, nodes, led, pods, gang_of, hint=hint)
    assert g.tobytes() == og.tobytes(), f"seed {seed}: gang decisions differ"
    assert eng.stats()["n_allocations"] == len(ogl), f"seed {seed}: ledger size after gangs differs"
    eng.close()
    return N, P, kind, flags


def permute_node(abi, node, new_of_old, n_new):
    """The same node with its GPU list re-ordered / shortened: the GPU at old index i moves to new_of_old[i] (-1: gone)."""
    out = np.zeros(1, dtype=abi.NODE_IMAGE)[0]
    for name in ("mig_free", "node_flags", "flags"):
        out[name] = node[name]
    out["n_gpus"] = n_new
    for i, a in enumerate(new_of_old):
        if a < 0:
            continue
        for name in ("pcie_bw", "free_bytes", "util", "numa", "pcie_switch", "arch"):
            out[name][a] = node[name][i]
        for bit_name in ("healthy_mask", "mig_enabled_mask", "mig_has_free_mask"):
            if int(node[bit_name]) >> i & 1:
                out[bit_name] |= 1 << a
        for j, b in enumerate(new_of_old):
            if b < 0:
                continue
            out["nvl_bw"][a][b] = node["nvl_bw"][i][j]
            if int(node["nvl_present"][i]) >> j & 1:
                out["nvl_present"][a] |= 1 << b
            if int(node["nvl_active"][i]) >> j & 1:
                out["nvl_active"][a] |= 1 << b
    return out


def life_cycle_seed(kg, orc, seed, n_ops=80):
    """Random interleaving of informer events (upsert / remove / re-add), ledger calls (commit, overwrite, release)
    and scoring / scheduling calls on ONE engine, mirrored on a plain model (node array + presence + ledger dict)
    that is handed to the oracle after every step."""
    abi, synth = kg.abi, kg.synth
    rng = synth.SplitMix64(seed * 104729 + 7)
    N = int(rng.randint(6, 90, 1)[0])
    nodes = synth.make_cluster(N, "mixed", seed=seed, unhealthy_frac=0.1, inactive_link_frac=0.05, partial_node_frac=0.3, mig_frac=0.3)
    pool = synth.make_cluster(4 * N, "mixed", seed=seed + 1, unhealthy_frac=0.2, inactive_link_frac=0.1, partial_node_frac=0.4, mig_frac=0.3)
    present = np.ones(N, dtype=bool)
    ledger = {}
    eng = kg.Engine(max_nodes=N + 3)
    eng.upsert_nodes(nodes)
    next_wl = [1000]

    def view():          # what the oracle sees: absent records have no GPUs
        v = nodes.copy()
        v["n_gpus"][~present] = 0; v["healthy_mask"][~present] = 0
        led = np.zeros(len(ledger), dtype=abi.ALLOC)
        for i, (wl, (n, m)) in enumerate(sorted(ledger.items())):
            led[i]["wl_id"], led[i]["node_idx"], led[i]["gpu_mask"] = wl, n, m
        return v, led

    def some_pods(n, k_lo):
        pods = synth.make_pods(n, seed=int(rng.randint(0, 1 << 30, 1)[0]), k_lo=k_lo, k_hi=8, pref_probs=(0.2, 0.3, 0.1, 0.2, 0.2), arch_frac=0.2, mig_frac=0.1)
        for i in range(n):       # mostly new ids, sometimes one that is already in the ledger (overwrite)
            if ledger and rng.uniform(1)[0] < 0.2:
                pods["wl_id"][i] = sorted(ledger)[int(rng.randint(0, len(ledger) - 1, 1)[0])]
            else:
                pods["wl_id"][i] = next_wl[0]; next_wl[0] += 1
        return pods

    for step in range(n_ops):
        op = int(rng.randint(0, 9, 1)[0])
        if op == 0 and rng.uniform(1)[0] < 0.35 and present.any():
            # a refresh that re-enumerates a node's GPUs (and sometimes loses one): the reference keys allocations by GPU UUID
            # (scheduler.go:586-598), the engine by index, so the refresh carries the old -> new map and the ledger moves with it
            i = int(np.flatnonzero(present)[int(rng.randint(0, int(present.sum()) - 1, 1)[0])])
            g = int(nodes["n_gpus"][i])
            if g >= 2:
                order = list(np.argsort(rng.uniform(g)))                     # new position of every old index
                lost = int(rng.randint(0, g - 1, 1)[0]) if rng.uniform(1)[0] < 0.4 else -1
                new_of_old, nxt = [-1] * 8, 0
                for old_idx in sorted(range(g), key=lambda q: order[q]):
                    if old_idx == lost:
                        continue
                    new_of_old[old_idx] = nxt; nxt += 1
                nodes[i] = permute_node(abi, nodes[i], new_of_old, nxt)
                eng.upsert_node_remap(i, nodes[i], new_of_old)
                for wl, (n2, m2) in list(ledger.items()):
                    if n2 == i:
                        m_new = 0
                        for b in range(8):
                            if m2 >> b & 1 and new_of_old[b] >= 0:
                                m_new |= 1 << new_of_old[b]
                        ledger[wl] = (i, m_new)                          # (an entry whose GPUs all vanished stays and still counts)
        elif op == 0:                                 # node refresh / re-add
            i = int(rng.randint(0, N - 1, 1)[0])
            nodes[i] = pool[int(rng.randint(0, len(pool) - 1, 1)[0])]
            eng.upsert_node(i, nodes[i]); present[i] = True
        elif op == 1:                                 # node delete (or an error for one that is gone)
            i = int(rng.randint(0, N - 1, 1)[0])
            if present[i]:
                eng.remove_node(i); present[i] = False
            else:
                try:
                    eng.remove_node(i); raise AssertionError(f"seed {seed}: removing an absent node succeeded")
                except kg.EngineError:
                    pass
        elif op == 2:                                 # external commit, sometimes over an existing id
            wl = sorted(ledger)[int(rng.randint(0, len(ledger) - 1, 1)[0])] if (ledger and rng.uniform(1)[0] < 0.3) else next_wl[0]
            next_wl[0] += 1
            i, m = int(rng.randint(0, N - 1, 1)[0]), int(rng.randint(1, 255, 1)[0])
            # what the engine must answer: a present node refuses GPUs it does not have and GPUs another workload holds
            others = 0
            for w2, (n2, m2) in ledger.items():
                if n2 == i and w2 != wl:
                    others |= m2
            bad = present[i] and ((m >> int(nodes["n_gpus"][i])) != 0 or (others & m) != 0)
            if bad:
                try:
                    eng.commit(wl, i, m); raise AssertionError(f"seed {seed}: a conflicting commit succeeded")
                except kg.EngineError as ex:
                    assert ex.code in (-1, -6)
            else:
                eng.commit(wl, i, m); ledger[wl] = (i, m)
        elif op == 3:                                 # release (known or unknown)
            if ledger and rng.uniform(1)[0] < 0.8:
                wl = sorted(ledger)[int(rng.randint(0, len(ledger) - 1, 1)[0])]
                eng.release(wl); del ledger[wl]
            else:
                try:
                    eng.release(999_999_999); raise AssertionError(f"seed {seed}: releasing an unknown workload succeeded")
                except kg.EngineError:
                    pass
        elif op in (4, 5):                            # a short queue with commits
            pods = some_pods(int(rng.randint(1, 6, 1)[0]), 1)
            v, led = view()
            d = eng.schedule_queue(pods)
            od, oled, _ = orc.schedule_queue(abi, v, led, pods)
            assert d.tobytes() == od.tobytes(), f"seed {seed} step {step}: queue decisions differ"
            ledger.clear()
            for a in oled:
                ledger[int(a["wl_id"])] = (int(a["node_idx"]), int(a["gpu_mask"]))
        else:                                         # stateless scoring (matrix + top-3, or the latency path)
            pods = some_pods(int(rng.randint(1, 12, 1)[0]), 0)
            v, led = view()
            if op == 6:
                _, t = eng.score_batch(pods, N, want_scores=False, topk=1)
                keep = pods["gpu_count"] >= 1        # a request for 0 GPUs would "fit" the oracle's empty stand-in for an absent node
                _, ot, _ = orc.score_batch(abi, v, led, pods, want_scores=False, topk=1)
                assert t[keep].tobytes() == ot[keep].tobytes(), f"seed {seed} step {step}: top-1 differs"
            else:
                sc, _ = eng.score_batch(pods, N, want_scores=True)
                o, _, _ = orc.score_batch(abi, v, led, pods, want_scores=True)
                assert sc[:, present].tobytes() == o[:, present].tobytes(), f"seed {seed} step {step}: score matrix differs"
                assert (sc[:, ~present]["flags"] == 0).all(), f"seed {seed} step {step}: an absent node was scored"
        assert eng.stats()["n_allocations"] == len(ledger), f"seed {seed} step {step}: ledger size"
    # checkpoint -> fresh engine answers the same
    blob = eng.snapshot()
    other = kg.Engine(max_nodes=N + 3)
    other.restore(blob)
    pods = some_pods(10, 0)
    a, _ = eng.score_batch(pods, N, want_scores=True)
    b, _ = other.score_batch(pods, N, want_scores=True)
    assert a.tobytes() == b.tobytes(), f"seed {seed}: restored engine differs"
    eng.close(); other.close()
    return N, len(ledger), int(present.sum())


def mig_seed(kg, orc, seed):
    """E2 on random tables: any occupancy byte (also ones no legal placement sequence produces), every profile id
    0..12 in the request stream, random target distributions; bin-packing and rebalancing vs the oracle."""
    abi, synth = kg.abi, kg.synth
    rng = synth.SplitMix64(seed * 15485863 + 3)
    G, R = int(rng.randint(1, 700, 1)[0]), int(rng.randint(1, 2500, 1)[0])
    gpus, profiles = synth.make_mig(G, R, seed=seed, occupancy=float(rng.uniform(1)[0]))
    if rng.uniform(1)[0] < 0.5:
        gpus["occ"] = rng.randint(0, 255, G).astype(np.uint8)
    if rng.uniform(1)[0] < 0.5:
        profiles = rng.randint(0, 12, R).astype(np.uint8)
    eng = kg.Engine(max_nodes=1)
    a, after = eng.mig_pack(gpus, profiles)
    o, o_after = orc.mig_pack(abi, gpus, profiles)
    assert a.tobytes() == o.tobytes() and after.tobytes() == o_after.tobytes(), f"seed {seed}: MIG packing differs"
    t = rng.uniform(10) * (rng.uniform(10) < 0.4)
    target = t / t.sum() * float(rng.uniform(1)[0]) if t.sum() > 0 else t
    for tbl in (gpus, after):
        ops, supply, n = eng.mig_rebalance(tbl, target)
        o_ops, o_supply, o_n = orc.mig_rebalance(abi, tbl, target)
        assert n == o_n and ops.tobytes() == o_ops.tobytes() and supply.tolist() == o_supply.tolist(), f"seed {seed}: MIG rebalance differs"
    eng.close()
    return G, R


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    first = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    kg = entry.load_package()
    from oracle import binding as orc
    for seed in range(first, first + n):
        N, P, kind, flags = one_seed(kg, orc, seed)
        print(f"seed {seed}: ok ({N} nodes, {P} pods, bandwidth kind {kind}, engine flags {flags})", flush=True)
        N, L, pres = life_cycle_seed(kg, orc, seed)
        print(f"seed {seed}: life cycle ok ({N} nodes, {pres} present at the end, {L} ledger entries)", flush=True)
        G, R = mig_seed(kg, orc, seed)
        print(f"seed {seed}: MIG ok ({G} GPUs, {R} requests)", flush=True)
    print(f"fuzz_parity: {n} seeds bit-exact")
