"""Seeded synthetic clusters and pod queues for the BASELINE.json configs.

Deterministic: a counter-based splitmix64 (seed = 0x4B475745 ^ config number) drives
every draw, so the oracle and the engine see byte-identical inputs on any machine.
Output is in the ABI layouts of include/kgwe_engine.h (numpy structured arrays).
"""
import numpy as np

from . import abi

GIB = 1 << 30
SEED_BASE = 0x4B475745
_M64 = (1 << 64) - 1


class SplitMix64:
    """Vectorised splitmix64 stream."""

    def __init__(self, seed):
        self.state = np.uint64(seed & _M64)

    def next_u64(self, n):
        with np.errstate(over="ignore"):
            idx = np.arange(1, n + 1, dtype=np.uint64)
            z = self.state + idx * np.uint64(0x9E3779B97F4A7C15)
            self.state = self.state + np.uint64(n) * np.uint64(0x9E3779B97F4A7C15)
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            return z ^ (z >> np.uint64(31))

    def uniform(self, n):
        return (self.next_u64(n) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))

    def randint(self, lo, hi, n):
        """Integers in [lo, hi]."""
        return (lo + (self.next_u64(n) % np.uint64(hi - lo + 1)).astype(np.int64)).astype(np.int64)

    def choice(self, probs, n):
        u = self.uniform(n)
        edges = np.cumsum(np.asarray(probs, dtype=np.float64))
        return np.minimum(np.searchsorted(edges, u, side="right"), len(probs) - 1)


_CUBE = np.zeros((8, 8), dtype=bool)          # 12-edge cube mesh of reference docs/architecture.md:275-286
for _i in range(8):
    for _b in (1, 2, 4):
        _CUBE[_i, _i ^ _b] = True
_FULL = ~np.eye(8, dtype=bool)


def _mask_bits(adj):
    """adj[..., i, j] bool -> uint8 per i with bit j."""
    w = (1 << np.arange(8)).astype(np.uint16)
    return (adj.astype(np.uint16) * w).sum(axis=-1).astype(np.uint8)


def make_cluster(n_nodes, kind="mixed", seed=SEED_BASE, unhealthy_frac=0.02, inactive_link_frac=0.01,
                 partial_node_frac=0.0, mig_frac=0.0):
    """kind: 'dgx' = all Hopper full mesh 900 GB/s (config 1); 'mixed' = 50% H100 / 30% A100
    (10% of those cube-mesh) / 20% B200 (configs 2, 3, 5)."""
    rng = SplitMix64(seed)
    N = n_nodes
    nodes = np.zeros(N, dtype=abi.NODE_IMAGE)
    if kind == "dgx":
        cls = np.zeros(N, dtype=np.int64)
    else:
        cls = rng.choice([0.5, 0.3, 0.2], N)             # 0 H100, 1 A100, 2 B200
    cube = (cls == 1) & (rng.uniform(N) < 0.10)
    bw = np.array([900.0, 600.0, 1800.0])[cls]
    arch = np.array([abi.ARCH_HOPPER, abi.ARCH_AMPERE, abi.ARCH_BLACKWELL], dtype=np.uint8)[cls]
    a100_small = (cls == 1) & (rng.uniform(N) < 0.5)
    total_gib = np.where(cls == 2, 180, np.where(a100_small, 40, 80)).astype(np.int64)
    n_gpus = np.full(N, 8, dtype=np.int64)
    if partial_node_frac > 0:
        part = rng.uniform(N) < partial_node_frac
        n_gpus = np.where(part, rng.randint(1, 7, N), n_gpus)
    gvalid = np.arange(8)[None, :] < n_gpus[:, None]                   # [N,8]
    adj = np.where(cube[:, None, None], _CUBE[None], _FULL[None])        # [N,8,8]
    adj = adj & gvalid[:, :, None] & gvalid[:, None, :]
    inactive = rng.uniform(N * 64).reshape(N, 8, 8) < inactive_link_frac
    inactive = inactive | inactive.transpose(0, 2, 1)                    # links go down in both directions
    nodes["nvl_bw"] = np.where(adj, bw[:, None, None], 0.0)
    nodes["nvl_present"] = _mask_bits(adj)
    nodes["nvl_active"] = _mask_bits(adj & ~inactive)
    nodes["pcie_bw"] = np.where(gvalid, np.where(cls == 1, 32.0, 64.0)[:, None], 0.0)
    free_gib = np.minimum(rng.randint(0, 40, N * 8).reshape(N, 8) + total_gib[:, None] // 2, total_gib[:, None])
    nodes["free_bytes"] = np.where(gvalid, free_gib.astype(np.uint64) * np.uint64(GIB), np.uint64(0))
    nodes["util"] = np.where(gvalid, np.floor(rng.uniform(N * 8).reshape(N, 8) * 200.0) / 2.0, 0.0)
    nodes["numa"] = (np.arange(8) // 4)[None, :].astype(np.int8)
    nodes["pcie_switch"] = (np.arange(8) // 2)[None, :].astype(np.uint8)
    nodes["arch"] = np.where(gvalid, arch[:, None], 0).astype(np.uint8)
    unhealthy = rng.uniform(N * 8).reshape(N, 8) < unhealthy_frac
    nodes["healthy_mask"] = _mask_bits(gvalid & ~unhealthy)
    nodes["n_gpus"] = n_gpus.astype(np.uint8)
    if mig_frac > 0:
        has = rng.uniform(N) < mig_frac
        cnt = rng.randint(0, 7, N * abi.MIG_PROFILES).reshape(N, abi.MIG_PROFILES)
        sparse = rng.uniform(N * abi.MIG_PROFILES).reshape(N, abi.MIG_PROFILES) < 0.4
        nodes["mig_free"] = np.where(has[:, None] & sparse, cnt, 0).astype(np.uint8)
        en = rng.uniform(N * 8).reshape(N, 8) < 0.3
        fr = rng.uniform(N * 8).reshape(N, 8) < 0.6
        nodes["mig_enabled_mask"] = _mask_bits(has[:, None] & en & gvalid)
        nodes["mig_has_free_mask"] = _mask_bits(has[:, None] & en & fr & gvalid)
    nodes["node_flags"] = (np.where(cube | (n_gpus < 8), 0, abi.NODE_NVSWITCH) | np.where(adj.any(axis=(1, 2)), abi.NODE_ANY_NVLINK, 0)).astype(np.uint8)
    nodes["flags"] = 1
    return nodes


def make_hint_reqs(n, seed=SEED_BASE):
    """discovery.GPURequirements for GetTopologyHint (discovery.go:250-281), the fields it reads."""
    rng = SplitMix64(seed ^ 0x48494E54)
    r = np.zeros(n, dtype=abi.HINT_REQ)
    r["count"] = rng.randint(0, 9, n)
    r["min_memory_gb"] = np.array([0, 16, 40, 80, 120])[rng.randint(0, 4, n)]
    r["prefer_nvlink"] = rng.randint(0, 1, n)
    r["prefer_same_numa"] = rng.randint(0, 1, n)
    r["arch"] = np.array([abi.ARCH_ANY, abi.ARCH_HOPPER, abi.ARCH_AMPERE, abi.ARCH_BLACKWELL], dtype=np.uint8)[rng.randint(0, 3, n)]
    return r


def make_pods(n_pods, seed=SEED_BASE, k_lo=1, k_hi=8, pref_probs=(0.2, 0.5, 0.0, 0.2, 0.1), arch_frac=0.0,
              minbw_choices=(0.0,), mig_frac=0.0, exhaustive=False, wl_base=1_000_000):
    """pref_probs indexes KGWE_TOPO_* (None, NVLinkOptimal, NVLinkRequired, SameNUMA, SamePCIeSwitch)."""
    rng = SplitMix64(seed ^ 0xA5A5A5A5)
    P = n_pods
    pods = np.zeros(P, dtype=abi.POD_REQ)
    pods["wl_id"] = wl_base + np.arange(P, dtype=np.uint64)
    pods["gpu_count"] = rng.randint(k_lo, k_hi, P)
    pods["min_memory_gb"] = np.array([16, 40, 80])[rng.randint(0, 2, P)]
    pods["topology_pref"] = rng.choice(list(pref_probs), P).astype(np.uint8)
    want_arch = rng.uniform(P) < arch_frac
    arch_pick = np.array([abi.ARCH_HOPPER, abi.ARCH_AMPERE, abi.ARCH_BLACKWELL], dtype=np.uint8)[rng.randint(0, 2, P)]
    pods["arch"] = np.where(want_arch, arch_pick, abi.ARCH_ANY)
    pods["min_bandwidth_gbps"] = np.asarray(minbw_choices, dtype=np.float64)[rng.randint(0, len(minbw_choices) - 1, P)]
    want_mig = rng.uniform(P) < mig_frac
    prof = rng.randint(1, abi.MIG_PROFILES, P).astype(np.uint8)
    pods["mig_profile"] = np.where(want_mig, prof, abi.MIG_NONE)
    pods["mig_count"] = np.where(want_mig, rng.randint(1, 3, P), 0)
    if exhaustive:
        pods["flags"] = abi.POD_EXHAUSTIVE
    return pods


def make_ledger(nodes, frac, seed=SEED_BASE, wl_base=1):
    """Pre-existing allocations: on `frac` of the nodes, 1-3 workloads holding disjoint GPU sets."""
    rng = SplitMix64(seed ^ 0x5EED1ED6)
    N = len(nodes)
    pick = np.nonzero(rng.uniform(N) < frac)[0]
    n_wl = rng.randint(1, 3, len(pick))
    sizes = rng.randint(1, 3, len(pick) * 3).reshape(len(pick), 3)
    out = []
    wl = wl_base
    for row, n in enumerate(pick):
        g = 0
        for j in range(int(n_wl[row])):
            s = int(sizes[row, j])
            if g + s > int(nodes["n_gpus"][n]):
                break
            mask = ((1 << s) - 1) << g
            out.append((wl, int(n), mask))
            wl += 1
            g += s
    led = np.zeros(len(out), dtype=abi.ALLOC)
    if out:
        led["wl_id"], led["node_idx"], led["gpu_mask"] = zip(*out)
    return led


def make_hint(n_pods, n_nodes, frac, seed=SEED_BASE):
    """PlacementHint.PreferredNodes as a bitmap [P, ceil(N/8)]."""
    rng = SplitMix64(seed ^ 0x41494E54)
    bits = rng.uniform(n_pods * n_nodes).reshape(n_pods, n_nodes) < frac
    return np.packbits(bits, axis=1, bitorder="little")


def make_mig(n_gpus, n_reqs, seed=SEED_BASE ^ 4, occupancy=0.4):
    """BASELINE configs[3]: MIG-enabled GPUs (half 40 GB-class, half 80 GB-class, 3% MIG-disabled) with
    a random VALID initial occupancy (~`occupancy` of the GPUs carry instances), and a request stream
    with the example ProfileDistribution 70/20/10 % for 1g/2g/3g (reference
    examples/distributed-training.yaml:183-186) plus 5 % 4g/7g.  Returns (gpus, profiles)."""
    rng = SplitMix64(seed)
    gpus = np.zeros(n_gpus, dtype=abi.MIG_GPU)
    gpus["cls"] = rng.randint(0, 1, n_gpus)
    gpus["enabled"] = (rng.uniform(n_gpus) >= 0.03).astype(np.uint8)
    # valid occupancies = unions of disjoint legal placements: pick a few random placements per busy GPU
    masks = np.array([1 << s for s in range(7)] + [3, 12, 48] + [15, 240] + [15] + [255], dtype=np.uint16)
    busy = rng.uniform(n_gpus) < occupancy
    occ = np.zeros(n_gpus, dtype=np.uint16)
    for _ in range(3):
        pick = masks[rng.randint(0, len(masks) - 1, n_gpus)]
        ok = busy & ((occ & pick) == 0) & (rng.uniform(n_gpus) < 0.7)
        occ = np.where(ok, occ | pick, occ)
    gpus["occ"] = occ.astype(np.uint8)
    geo = rng.choice([0.665, 0.19, 0.095, 0.025, 0.025], n_reqs)      # 70/20/10 scaled to 95 % + 2.5 % 4g + 2.5 % 7g
    cls = rng.randint(0, 1, n_reqs)
    profiles = (2 * geo + 1 + cls).astype(np.uint8)
    return gpus, profiles


def config(num, scale=1.0):
    """(nodes, pods, ledger) for BASELINE.json configs[num-1]; `scale` shrinks N and P for tests."""
    seed = SEED_BASE ^ num
    if num == 1:    # 16-node DGX, 100 pending 1-8 GPU pods
        nodes = make_cluster(16, "dgx", seed, inactive_link_frac=0.0)
        pods = make_pods(100, seed)
        return nodes, pods, make_ledger(nodes, 0.0, seed)
    if num == 2:    # 1,000-node mixed cluster, 1k pods, NVLinkOptimal + minBandwidth filter
        n, p = max(8, int(1000 * scale)), max(8, int(1000 * scale))
        nodes = make_cluster(n, "mixed", seed)
        pods = make_pods(p, seed, k_lo=2, pref_probs=(0, 1, 0, 0, 0), arch_frac=0.3, minbw_choices=(0.0, 300.0, 600.0, 900.0))
        return nodes, pods, make_ledger(nodes, 0.3, seed)
    if num == 3:    # 10,000-node x 8-GPU cluster, 10k-pod burst, subset scoring
        n, p = max(8, int(10000 * scale)), max(8, int(10000 * scale))
        nodes = make_cluster(n, "mixed", seed)
        pods = make_pods(p, seed, pref_probs=(0, 1, 0, 0, 0))
        return nodes, pods, make_ledger(nodes, 0.3, seed)
    if num == 5:    # gang-scheduled worldSize=64: 8 member pods x k=8 per gang, NUMA/PCIe/NVLink mix
        n, gangs = max(8, int(10000 * scale)), max(1, int(1250 * scale))
        nodes = make_cluster(n, "mixed", seed)
        pods = make_pods(gangs * 8, seed, k_lo=8, k_hi=8, pref_probs=(0, 0.6, 0, 0.2, 0.2))
        return nodes, pods, make_ledger(nodes, 0.1, seed)
    raise ValueError(f"config {num} is not a scoring config")
