"""CPU checks of the extension oracles (own spec, no reference counterpart): hand-derived
values for E4 side terms and E5 gang roll-back."""
import numpy as np

from helpers import CUBE_EDGES, ledger_of, make_node, make_pod


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
