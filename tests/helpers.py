"""Hand-built node images / pods for known-answer tests (not synthetic-random)."""
import numpy as np

GIB = 1 << 30
CUBE_EDGES = [(0, 1), (0, 2), (0, 4), (1, 3), (1, 5), (2, 3), (2, 6), (3, 7), (4, 5), (4, 6), (5, 7), (6, 7)]


def make_node(abi, n_gpus=8, links="full", bw=900.0, free_gib=80, util=0.0, arch=None, numa=None, switch=None,
              unhealthy=(), inactive=(), pcie_bw=64.0, mig_free=None, bw_matrix=None):
    """links: 'full', 'none', or a list of undirected (i, j) edges.  inactive: edges present but IsActive=false."""
    arch = abi.ARCH_HOPPER if arch is None else arch
    n = np.zeros(1, dtype=abi.NODE_IMAGE)[0]
    if links == "full":
        edges = [(i, j) for i in range(n_gpus) for j in range(i + 1, n_gpus)]
    elif links == "none":
        edges = []
    else:
        edges = list(links)
    for (i, j) in edges:
        for a, b in ((i, j), (j, i)):
            n["nvl_bw"][a][b] = bw
            n["nvl_present"][a] |= 1 << b
            if (i, j) not in inactive and (j, i) not in inactive:
                n["nvl_active"][a] |= 1 << b
    if bw_matrix is not None:          # directed overrides: bw_matrix[i][j] = GPU i's entry for GPU j (None keeps the default)
        for i, row in enumerate(bw_matrix):
            for j, v in enumerate(row):
                if v is not None and i != j:
                    n["nvl_bw"][i][j] = v
    for g in range(n_gpus):
        n["pcie_bw"][g] = pcie_bw
        n["free_bytes"][g] = (free_gib[g] if isinstance(free_gib, (list, tuple)) else free_gib) * GIB
        n["util"][g] = util[g] if isinstance(util, (list, tuple)) else util
        n["arch"][g] = arch[g] if isinstance(arch, (list, tuple)) else arch
        n["numa"][g] = numa[g] if numa is not None else g // 4
        n["pcie_switch"][g] = switch[g] if switch is not None else g // 2
    n["n_gpus"] = n_gpus
    n["healthy_mask"] = ((1 << n_gpus) - 1) & ~sum(1 << g for g in unhealthy)
    if mig_free:
        for p, c in mig_free.items():
            n["mig_free"][p - 1] = c
    n["flags"] = 1
    return n


def make_pod(abi, k, pref="NVLinkOptimal", mem=40, arch=0, wl_id=1, mig=None, minbw=0.0, flags=0):
    prefs = {"None": abi.TOPO_NONE, "NVLinkOptimal": abi.TOPO_NVLINK_OPTIMAL, "NVLinkRequired": abi.TOPO_NVLINK_REQUIRED,
             "SameNUMA": abi.TOPO_SAME_NUMA, "SamePCIeSwitch": abi.TOPO_SAME_PCIE_SWITCH}
    p = np.zeros(1, dtype=abi.POD_REQ)[0]
    p["wl_id"] = wl_id
    p["gpu_count"] = k
    p["min_memory_gb"] = mem
    p["topology_pref"] = prefs[pref]
    p["arch"] = arch
    p["min_bandwidth_gbps"] = minbw
    p["flags"] = flags
    if mig:
        p["mig_profile"], p["mig_count"] = mig
    return p


def ledger_of(abi, entries):
    led = np.zeros(len(entries), dtype=abi.ALLOC)
    for i, (wl, node, mask) in enumerate(entries):
        led[i]["wl_id"], led[i]["node_idx"], led[i]["gpu_mask"] = wl, node, mask
    return led


def order_list(order):
    n = (int(order) >> 24) & 15
    return [(int(order) >> (3 * p)) & 7 for p in range(n)]
