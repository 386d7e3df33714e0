"""CPU checks of the extension oracles (own spec, no reference counterpart): hand-derived
values for E4 side terms and E5 gang roll-back."""
import numpy as np

from helpers import CUBE_EDGES, ledger_of, make_node, make_pod, order_list


def _pl(abi, node, mask):
    p = np.zeros(1, dtype=abi.PLACEMENT)[0]
    p["node_idx"], p["gpu_mask"] = node, mask
    return p


def test_side_terms_known_answers(kg, orc):
    abi = kg.abi
    nodes = np.array([make_node(abi), make_node(abi, links=CUBE_EDGES), make_node(abi, inactive=[(0, 1)]),
                      make_node(abi, n_gpus=4, links=[(0, 1), (2, 3)], bw=600.0)], dtype=abi.NODE_IMAGE)
    pls = np.array([
        _pl(abi, 0, 0xFF),   # full mesh 900, k=8: balanced cut has 4*4 edges -> 14400; every ring's weakest edge 900
        _pl(abi, 0, 0x0F),   # k=4: 2*2*900 = 3600; 900
        _pl(abi, 0, 0x07),   # k=3: split 1|2 -> 2 edges -> 1800; 900
        _pl(abi, 0, 0x03),   # k=2: 900; 900
        _pl(abi, 0, 0x01),   # k=1: 0; 0
        _pl(abi, 1, 0xFF),   # 3-cube: minimum bisection is 4 edges = 3600; the cube has a Hamiltonian cycle -> 900
        _pl(abi, 1, 0x07),   # GPUs 0,1,2 of the cube: edges 0-1, 0-2 only; best 1|2 split cuts one edge -> 900; no triangle -> ring 0
        _pl(abi, 2, 0x03),   # the pair whose link is inactive: 0; 0
        _pl(abi, 3, 0x0F),   # two disjoint pairs: {0,1}|{2,3} cuts nothing -> 0; no ring -> 0
        _pl(abi, abi.NO_NODE, 0),
    ], dtype=abi.PLACEMENT)
    t = orc.placement_terms(abi, nodes, pls)
    want = [(14400, 900), (3600, 900), (1800, 900), (900, 900), (0, 0), (3600, 900), (900, 0), (0, 0), (0, 0), (0, 0)]
    assert [(float(a), float(b)) for a, b in zip(t["bisection_gbps"], t["ring_bottleneck_gbps"])] == [(float(a), float(b)) for a, b in want]


def test_gang_rollback(kg, orc):
    abi = kg.abi
    nodes = np.array([make_node(abi), make_node(abi)], dtype=abi.NODE_IMAGE)
    # gang 0: two 4-GPU members -> node 0 (100), then node 1 (100 beats node 0's 96).
    # gang 1: three 4-GPU members -> node 0, node 1, then nothing is left -> the whole gang is rolled back.
    # gang 2: one 4-GPU member -> must see the GPUs gang 1 gave back: node 0, GPUs 4-7.
    pods = np.array([make_pod(abi, 4, wl_id=1), make_pod(abi, 4, wl_id=2),
                     make_pod(abi, 4, wl_id=3), make_pod(abi, 4, wl_id=4), make_pod(abi, 4, wl_id=5),
                     make_pod(abi, 4, wl_id=6)], dtype=abi.POD_REQ)
    gang_of = np.array([0, 0, 1, 1, 1, 2], dtype=np.uint32)
    for mode in (0, 1):
        dec, led = orc.schedule_gangs(abi, nodes, None, pods, gang_of, mode=mode)
        assert dec[0]["node_idx"] == 0 and dec[1]["node_idx"] == 1
        assert (dec[2:5]["node_idx"] == abi.NO_NODE).all()
        assert [int(f) for f in dec[2:5]["flags"]] == [abi.F_GANG_ROLLBACK, abi.F_GANG_ROLLBACK, 0]
        assert dec[5]["node_idx"] == 0 and dec[5]["gpu_mask"] == 0xF0 and len(led) == 3


def test_mig_pack_known_answers(kg, orc):
    """Hand-derived best-fit cases on one 80 GB-class GPU (profile ids: 2=1g.10gb 4=2g.20gb 6=3g.40gb 8=4g.40gb 10=7g.80gb)."""
    abi = kg.abi
    def gpus(*occ, cls=1, enabled=1):
        g = np.zeros(len(occ), dtype=abi.MIG_GPU)
        g["occ"], g["cls"], g["enabled"] = occ, cls, enabled
        return g
    # empty GPU, 1g: start s overlaps {1g@s, the 2g covering s (none for s=6), the 3g covering s, 4g if s<4, 7g}:
    #   s=0..3 -> 5, s=4,5 -> 4, s=6 -> 3  => best-fit takes start 6 (cost 3)
    a, g = orc.mig_pack(abi, gpus(0), [2])
    assert (a[0]["gpu_idx"], a[0]["start"], a[0]["size"], a[0]["cost"]) == (0, 6, 1, 3) and g[0]["occ"] == 0x40
    # then a 3g: start 0 overlaps 1g x4 + 2g x2 + 3g@0 + 4g + (7g already dead) = 8; start 4 is blocked by slice 6 -> start 0
    a, g = orc.mig_pack(abi, g, [6])
    assert (a[0]["start"], a[0]["cost"]) == (0, 8) and g[0]["occ"] == 0x4F
    # 4g no longer fits (slices 0-3 taken), 7g neither; 2g fits only at start 4 (1g@4, 1g@5, 2g@4 -> cost 3)
    a, g = orc.mig_pack(abi, g, [8, 10, 4])
    assert a[0]["gpu_idx"] == abi.NO_NODE and a[1]["gpu_idx"] == abi.NO_NODE
    assert (a[2]["gpu_idx"], a[2]["start"], a[2]["cost"]) == (0, 4, 3) and g[0]["occ"] == 0x7F
    # class and enable filters; ties -> lowest GPU index; the fuller GPU wins when it destroys less
    a, g = orc.mig_pack(abi, gpus(0, 0, cls=[0, 1]), [1, 2, 3])          # 1g.5gb -> GPU0, 1g.10gb -> GPU1, 2g.10gb -> GPU0
    assert a["gpu_idx"].tolist() == [0, 1, 0]
    a, g = orc.mig_pack(abi, gpus(0, 0, enabled=[0, 1]), [2])
    assert a[0]["gpu_idx"] == 1
    a, g = orc.mig_pack(abi, gpus(0x00, 0x0F), [4])     # empty GPU: cheapest 2g is start 4 (cost 5); GPU with 0-3 taken: start 4 costs 4 -> GPU 1
    assert (a[0]["gpu_idx"], a[0]["start"], a[0]["cost"]) == (1, 4, 4)
    a, _ = orc.mig_pack(abi, gpus(0), [0, 11])          # invalid profile ids never fit
    assert (a["gpu_idx"] == abi.NO_NODE).all()


def test_mig_rebalance_known_answers(kg, orc):
    abi = kg.abi
    g = np.zeros(12, dtype=abi.MIG_GPU)
    g["enabled"] = 1
    g["cls"] = [1] * 10 + [0] * 2
    g["occ"][3] = 1                                     # busy -> not idle; 9 idle 80 GB GPUs, 2 idle 40 GB GPUs
    target = np.zeros(10)
    target[1], target[3], target[5] = 0.7, 0.2, 0.1    # 1g.10gb / 2g.20gb / 3g.40gb (the example strategy)
    ops, supply, n = orc.mig_rebalance(abi, g, target)
    # shares of 9: 6.3 / 1.8 / 0.9 -> floors 6,1,0; the 2 left go to the largest fractions .9 (3g) then .8 (2g) -> 6,2,1
    assert supply.tolist() == [0, 42, 0, 6, 0, 2, 0, 0, 0, 0] and n == 50 and len(ops) == 50
    assert ops["gpu_idx"][:7].tolist() == [0] * 7 and ops["start"][:7].tolist() == list(range(7))
    assert ops["gpu_idx"][42:48].tolist() == [7, 7, 7, 8, 8, 8] and ops["start"][42:48].tolist() == [0, 2, 4, 0, 2, 4]   # idle order skips GPU 3
    assert ops["gpu_idx"][48:].tolist() == [9, 9] and ops["start"][48:].tolist() == [0, 4] and ops["profile"][48] == 6
    # the 40 GB class has no target -> untouched
    assert not (ops["gpu_idx"] >= 10).any()


def test_mig_reconfigure_known_answers(kg, orc):
    """The reconfiguration pass with destroy ops, worked by hand.  Placement bits: 0..6 = 1g at slice 0..6, 7..9 = 2g at
    0/2/4, 10..11 = 3g at 0/4, 12 = 4g, 13 = 7g.  Five 80 GB-class GPUs (class 1), target 1g.10gb : 3g.40gb = 1 : 1."""
    abi = kg.abi
    g = np.zeros(5, dtype=abi.MIG_GPU_STATE)
    g["cls"], g["enabled"] = 1, 1
    g["created"][0] = 0b0000000_1111111                 # GPU 0: seven 1g instances, none in use - already what 1g wants
    g["created"][1] = (1 << 7) | (1 << 10)               # GPU 1: a 2g at 0 and a 3g at 0 (an impossible mix on real hardware; the pass only reads bits)
    g["created"][2] = 1 << 13; g["in_use"][2] = 1 << 13   # GPU 2: its 7g instance is running: left alone
    g["created"][3] = 0                                  # GPU 3: nothing yet
    g["created"][4] = (1 << 11) | (1 << 0) | (1 << 8); g["in_use"][4] = 1 << 0      # GPU 4: a 1g running -> left alone; its free 3g and 2g count as supply
    target = np.zeros(10); target[1] = 0.5; target[5] = 0.5
    ops, supply, n = orc.mig_reconfigure(abi, g, target)
    # reconfigurable: GPUs 0, 1, 3 -> shares 1.5 / 1.5: one each, the third to the lower profile (tie) -> 1g gets GPUs 0, 1; 3g gets GPU 3
    want = []
    # GPU 0 (1g): nothing to do.  GPU 1 (1g): destroy 2g@0 (bit 7) and 3g@0 (bit 10), create 1g at 0..6
    want += [(1, 4, 0, 2, 1), (1, 6, 0, 4, 1)] + [(1, 2, s, 1, 0) for s in range(7)]
    # GPU 3 (3g): create 3g at 0 and 4
    want += [(3, 6, 0, 4, 0), (3, 6, 4, 4, 0)]
    got = [(int(o["gpu_idx"]), int(o["profile"]), int(o["start"]), int(o["size"]), int(o["pad"])) for o in ops]
    assert n == len(want) and got == want
    # supply: 1g.10gb: 2 GPUs x 7; 3g.40gb: GPU 3's two + GPU 4's free one; 2g.20gb: GPU 4's free one
    assert supply.tolist() == [0, 14, 0, 1, 0, 3, 0, 0, 0, 0]


def test_topology_hint_known_answers(kg, orc):
    """GetTopologyHint (discovery.go:222-247, 378-488), hand-derived: base 50, +30 linked group, +10 same NUMA,
    +5 per selected GPU of the wanted architecture; availability = healthy and (free MIG instance | util < 90)."""
    abi = kg.abi
    busy = make_node(abi, util=[95.0] * 6 + [10.0, 10.0])                 # only GPUs 6,7 are under 90 %
    cube = make_node(abi, links=CUBE_EDGES, arch=abi.ARCH_AMPERE)
    small = make_node(abi, free_gib=20)
    mig = make_node(abi, util=99.0)
    mig["mig_enabled_mask"], mig["mig_has_free_mask"] = 0x0F, 0x05        # GPUs 0 and 2 have a free instance; 4-7 fail on util
    nodes = np.array([busy, cube, small, mig], dtype=abi.NODE_IMAGE)
    def req(count, mem=0, nvl=0, numa=0, arch=0):
        r = np.zeros(1, dtype=abi.HINT_REQ)
        r["count"], r["min_memory_gb"], r["prefer_nvlink"], r["prefer_same_numa"], r["arch"] = count, mem, nvl, numa, arch
        return r
    def ask(r, nn=nodes):
        t = orc.topology_hints(abi, nn, r)[0]
        return int(t["node_idx"]) if t["node_idx"] != abi.NO_NODE else None, int(t["score"]), order_list(t["order"]), float(t["bandwidth_estimate"])
    # 2 GPUs with NVLink + same NUMA + Hopper: node 0 -> GPUs 6,7 linked, NUMA 1, both Hopper: 50+30+10+10 = 100
    assert ask(req(2, 40, 1, 1, abi.ARCH_HOPPER)) == (0, 100, [6, 7], 900.0)
    # 4 GPUs: node 0 has only 2 available; cube node: GPU 0 has 3 linked peers (1,2,4) -> [0,1,2,4], NUMA differs; Ampere wanted: 50+30+20 = 100
    assert ask(req(4, 40, 1, 1, abi.ARCH_AMPERE)) == (1, 100, [0, 1, 2, 4], (900 * 3 + 64 * 3) / 6)
    # 5 GPUs on the cube: no GPU has 4 peers -> first 5, no NVLink bonus; node 2 (full mesh, 20 GiB free) wins with 80 when memory allows
    assert ask(req(5, 16, 1, 0, 0)) == (2, 80, [0, 1, 2, 3, 4], 900.0)
    assert ask(req(5, 40, 1, 0, 0)) == (1, 50, [0, 1, 2, 3, 4], (900 * 5 + 64 * 5) / 10)     # cube edges among 0..4: 0-1 0-2 0-4 1-3 2-3
    # MIG rule: node 3's only available GPUs are 0 and 2 (free instances), util does not matter for them
    assert ask(req(2, 0, 0, 1, 0), nodes[3:]) == (0, 60, [0, 2], 900.0)
    assert ask(req(3, 0, 0, 0, 0), nodes[3:]) == (None, 0, [], 0.0)       # "no suitable placement found for requirements"
    # count 1 / 0: no group search; count 0 selects nothing and still scores 50 (+10: an empty set is "same NUMA")
    assert ask(req(1, 0, 1, 0, 0)) == (0, 50, [6], 0.0)
    assert ask(req(0, 0, 1, 1, 0)) == (0, 60, [], 0.0)
    # exporter gauge: 50 base, +30 NVSwitch, +20 any NVLink entry, 0 without GPUs
    g = np.array([make_node(abi), make_node(abi, links="none"), make_node(abi, n_gpus=0)], dtype=abi.NODE_IMAGE)
    g["node_flags"] = [abi.NODE_NVSWITCH | abi.NODE_ANY_NVLINK, 0, abi.NODE_NVSWITCH]
    assert orc.topology_gauges(abi, g).tolist() == [100, 50, 0]
