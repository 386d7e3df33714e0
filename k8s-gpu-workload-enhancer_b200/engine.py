"""Thin ctypes wrapper over the C ABI (include/kgwe_engine.h).  One method per entry
point, numpy arrays in the ABI's layouts in and out.  Mirrors what the cgo shim in the
reference's src/scheduler package does; no scoring logic here."""
import ctypes as C

import numpy as np

from . import abi


class EngineError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"kgwe error {code}: {msg}")
        self.code = code


class Engine:
    """Owns one kgwe_engine (one CUDA device, one node shard)."""

    def __init__(self, max_nodes, device=0, node_base=0, config=None, flags=0):
        self.lib = abi.load_engine_lib()
        self.cfg = config if config is not None else abi.default_config(max_nodes, device, node_base)
        self.cfg.flags |= flags
        self.h = C.c_void_p()
        rc = self.lib.kgwe_engine_create(C.byref(self.cfg), C.byref(self.h))
        if rc != 0:
            raise EngineError(rc, "kgwe_engine_create failed (no usable CUDA device? there is no CPU fallback)")

    def close(self):
        if self.h:
            self.lib.kgwe_engine_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            msg = self.lib.kgwe_last_error(self.h)
            raise EngineError(rc, msg.decode() if msg else "")

    # --- snapshot ingestion (replaces discovery.go:290-319 full rebuilds) ---
    def upsert_nodes(self, images, first=0):
        images = np.ascontiguousarray(images, dtype=abi.NODE_IMAGE)
        self._check(self.lib.kgwe_nodes_upsert(self.h, first, len(images), abi.ptr(images)))

    def upsert_node(self, idx, image):
        img = np.ascontiguousarray(image, dtype=abi.NODE_IMAGE).reshape(1)
        self._check(self.lib.kgwe_node_upsert(self.h, idx, abi.ptr(img)))

    def upsert_node_remap(self, idx, image, old_to_new):
        """Refresh of a node whose GPU list changed: old_to_new[i] = new index of the GPU formerly at index i, -1 if gone."""
        img = np.ascontiguousarray(image, dtype=abi.NODE_IMAGE).reshape(1)
        m = np.full(abi.MAX_GPUS, -1, dtype=np.int8)
        m[: len(old_to_new)] = old_to_new
        self._check(self.lib.kgwe_node_upsert_remap(self.h, idx, abi.ptr(img), abi.ptr(m)))

    def remove_node(self, idx):
        self._check(self.lib.kgwe_node_remove(self.h, idx))

    def read_node(self, idx):
        out = np.zeros(1, dtype=abi.NODE_IMAGE)
        self._check(self.lib.kgwe_node_read(self.h, idx, abi.ptr(out)))
        return out[0]

    # --- ledger (scheduler.go:695-727) ---
    def commit(self, wl_id, node_idx, gpu_mask):
        self._check(self.lib.kgwe_alloc_commit(self.h, int(wl_id), int(node_idx), int(gpu_mask)))

    def release(self, wl_id):
        self._check(self.lib.kgwe_alloc_release(self.h, int(wl_id)))

    def seed_ledger(self, ledger):
        for a in ledger:
            self.commit(a["wl_id"], a["node_idx"], a["gpu_mask"])

    # --- scoring ---
    def score_batch(self, pods, n_nodes, hint=None, want_scores=True, topk=0):
        """scoreNodes for every pod (scheduler.go:182-203).  Returns (scores[P,N] | None, topk[P,topk] | None)."""
        pods = np.ascontiguousarray(pods, dtype=abi.POD_REQ)
        P = len(pods)
        scores = np.zeros((P, n_nodes), dtype=abi.NODE_SCORE) if want_scores else None
        top = np.zeros((P, topk), dtype=abi.PLACEMENT) if topk else None
        if hint is not None:
            hint = np.ascontiguousarray(hint, dtype=np.uint8)
        self._check(self.lib.kgwe_score_batch(self.h, abi.ptr(pods), P, abi.ptr(hint), abi.ptr(scores), topk, abi.ptr(top)))
        return scores, top

    def score_batch_device(self, d_pods, n_pods, d_hint, d_scores, topk, d_topk, stream=None):
        """All arguments are raw device addresses (ints) on this engine's GPU; asynchronous."""
        self._check(self.lib.kgwe_score_batch_device(self.h, d_pods, n_pods, d_hint, d_scores, topk, d_topk, stream))

    def score_top1_noalloc(self, d_pods, n_pods, d_hint, d_scores, d_top1, stream=None):
        """kgwe_score_top1_noalloc: best placement per pod with the key table in the library's
        fixed device array (no allocation inside the call; at most abi.NOALLOC_MAX_PODS pods)."""
        self._check(self.lib.kgwe_score_top1_noalloc(self.h, d_pods, n_pods, d_hint, d_scores, d_top1, stream))

    def score_batch_device_p2p(self, d_pods, n_pods, d_hint, d_scores, peer_ptrs, my_rank, stream=None):
        """Top-1 scoring whose finalisation kernel writes this rank's placements into every rank's gather
        buffer (peer_ptrs: one device address per rank) over NVLink, replacing the all-gather."""
        arr = (C.c_uint64 * len(peer_ptrs))(*[int(x) for x in peer_ptrs])
        self._check(self.lib.kgwe_score_batch_device_p2p(self.h, d_pods, n_pods, d_hint, d_scores, arr, len(peer_ptrs), my_rank, stream))

    def score_merge_p2p(self, d_pods, n_pods, d_hint, d_scores, peer_ptrs, flag_ptrs, my_rank, epoch, d_merged, stream=None):
        """The whole sharded top-1 step: K1, finalise + P2P scatter + flag, wait + merge (kgwe_score_merge_p2p)."""
        a = (C.c_uint64 * len(peer_ptrs))(*[int(x) for x in peer_ptrs])
        f = (C.c_uint64 * len(flag_ptrs))(*[int(x) for x in flag_ptrs])
        self._check(self.lib.kgwe_score_merge_p2p(self.h, d_pods, n_pods, d_hint, d_scores, a, f, len(peer_ptrs), my_rank, epoch, d_merged, stream))

    def topk_merge_device(self, d_lists, n_lists, n_pods, topk, d_out, stream=None):
        self._check(self.lib.kgwe_topk_merge_device(self.h, d_lists, n_lists, n_pods, topk, d_out, stream))

    def schedule_queue(self, pods, hint=None):
        """Schedule() per pod in order with commits (scheduler.go:114-179)."""
        pods = np.ascontiguousarray(pods, dtype=abi.POD_REQ)
        out = np.zeros(len(pods), dtype=abi.PLACEMENT)
        if hint is not None:
            hint = np.ascontiguousarray(hint, dtype=np.uint8)
        self._check(self.lib.kgwe_schedule_queue(self.h, abi.ptr(pods), len(pods), abi.ptr(hint), abi.ptr(out)))
        return out

    def schedule_queue_p2p(self, pods, recv_ptrs, my_rank, seq_base, hint=None):
        """The stateful queue over a node-sharded cluster, exchanged on the device (kgwe_schedule_queue_p2p): every rank calls this
        with the same pods; the result holds the decisions whose node this rank owns, NO_NODE elsewhere."""
        pods = np.ascontiguousarray(pods, dtype=abi.POD_REQ)
        out = np.zeros(len(pods), dtype=abi.PLACEMENT)
        if hint is not None:
            hint = np.ascontiguousarray(hint, dtype=np.uint8)
        a = (C.c_uint64 * len(recv_ptrs))(*[int(x) for x in recv_ptrs])
        self._check(self.lib.kgwe_schedule_queue_p2p(self.h, abi.ptr(pods), len(pods), abi.ptr(hint), a, len(recv_ptrs), my_rank, seq_base, abi.ptr(out)))
        return out

    def schedule_gangs(self, pods, gang_of, hint=None):
        """E5: all-or-nothing gangs (members consecutive, gang_of non-decreasing)."""
        pods = np.ascontiguousarray(pods, dtype=abi.POD_REQ)
        gang_of = np.ascontiguousarray(gang_of, dtype=np.uint32)
        out = np.zeros(len(pods), dtype=abi.PLACEMENT)
        if hint is not None:
            hint = np.ascontiguousarray(hint, dtype=np.uint8)
        self._check(self.lib.kgwe_schedule_gangs(self.h, abi.ptr(pods), len(pods), abi.ptr(gang_of), abi.ptr(hint), abi.ptr(out)))
        return out

    def placement_terms(self, placements):
        """E4: bisection bandwidth and ring bottleneck of each placement's GPU set."""
        pl = np.ascontiguousarray(placements, dtype=abi.PLACEMENT).reshape(-1)
        out = np.zeros(len(pl), dtype=abi.SIDE_TERMS)
        self._check(self.lib.kgwe_placement_terms(self.h, abi.ptr(pl), len(pl), abi.ptr(out)))
        return out

    def mig_pack(self, gpus, profiles):
        """E2: serve MIG slice requests in order (best-fit).  Returns (assignments, gpus after)."""
        gpus = np.ascontiguousarray(gpus, dtype=abi.MIG_GPU).copy()
        profiles = np.ascontiguousarray(profiles, dtype=np.uint8)
        out = np.zeros(len(profiles), dtype=abi.MIG_ASSIGNMENT)
        self._check(self.lib.kgwe_mig_pack(self.h, abi.ptr(gpus), len(gpus), abi.ptr(profiles), len(profiles), abi.ptr(out)))
        return out, gpus

    def mig_rebalance(self, gpus, target, ops_cap=None):
        """E2: repartition idle GPUs toward `target` (10 fractions).  Returns (ops, supply[10])."""
        gpus = np.ascontiguousarray(gpus, dtype=abi.MIG_GPU)
        target = np.ascontiguousarray(target, dtype=np.float64)
        cap = ops_cap if ops_cap is not None else 7 * len(gpus)
        ops = np.zeros(cap, dtype=abi.MIG_OP)
        supply = np.zeros(10, dtype=np.uint32)
        n = C.c_uint32(0)
        self._check(self.lib.kgwe_mig_rebalance(self.h, abi.ptr(gpus), len(gpus), abi.ptr(target), abi.ptr(ops), cap, C.byref(n), abi.ptr(supply)))
        return ops[: min(n.value, cap)].copy(), supply, n.value

    def mig_reconfigure(self, gpus, target, ops_cap=None):
        """E2: reconfigure GPUs with nothing in use toward `target`, with destroy and create ops.  Returns (ops, supply[10], n_ops)."""
        gpus = np.ascontiguousarray(gpus, dtype=abi.MIG_GPU_STATE)
        target = np.ascontiguousarray(target, dtype=np.float64)
        cap = ops_cap if ops_cap is not None else 21 * len(gpus)
        ops = np.zeros(cap, dtype=abi.MIG_OP)
        supply = np.zeros(10, dtype=np.uint32)
        n = C.c_uint32(0)
        self._check(self.lib.kgwe_mig_reconfigure(self.h, abi.ptr(gpus), len(gpus), abi.ptr(target), abi.ptr(ops), cap, C.byref(n), abi.ptr(supply)))
        return ops[: min(n.value, cap)].copy(), supply, n.value

    def topology_hints(self, reqs):
        """DiscoveryService.GetTopologyHint for each request (discovery.go:222-247)."""
        reqs = np.ascontiguousarray(reqs, dtype=abi.HINT_REQ)
        out = np.zeros(len(reqs), dtype=abi.TOPOLOGY_HINT)
        self._check(self.lib.kgwe_topology_hint_batch(self.h, abi.ptr(reqs), len(reqs), abi.ptr(out)))
        return out

    def topology_gauges(self, n_nodes):
        """PrometheusExporter.calculateTopologyScore per node (prometheus_exporter.go:517-539)."""
        out = np.zeros(n_nodes, dtype=np.int32)
        self._check(self.lib.kgwe_topology_gauges(self.h, abi.ptr(out), n_nodes))
        return out

    def snapshot(self):
        """Checkpoint: cluster image + ledger as an opaque blob (bytes)."""
        n = C.c_uint64(0)
        self._check(self.lib.kgwe_snapshot_size(self.h, C.byref(n)))
        buf = np.zeros(n.value, dtype=np.uint8)
        w = C.c_uint64(0)
        self._check(self.lib.kgwe_snapshot_save(self.h, abi.ptr(buf), n.value, C.byref(w)))
        return buf[: w.value].tobytes()

    def restore(self, blob):
        buf = np.frombuffer(blob, dtype=np.uint8).copy()
        self._check(self.lib.kgwe_snapshot_load(self.h, abi.ptr(buf), len(buf)))

    def stats(self):
        s = abi.Stats()
        self._check(self.lib.kgwe_get_stats(self.h, C.byref(s)))
        d = {name: getattr(s, name) for name, _ in abi.Stats._fields_}
        d["latency_us_bucket"] = list(d["latency_us_bucket"])
        return d

    def latency_quantile(self, q):
        """Upper bucket edge (microseconds) below which a fraction q of the observed calls fell - what
        histogram_quantile() of the Grafana p99 panel reads off the same buckets."""
        st = self.stats()
        need, seen = q * st["latency_count"], 0
        for edge, n in zip(abi.LATENCY_BUCKET_EDGES_US, st["latency_us_bucket"]):
            seen += n
            if seen >= need and st["latency_count"]:
                return edge
        return None


def unpack_order(order):
    """SelectedGPUs (ordered) from the packed `order` field."""
    n = (int(order) >> 24) & 15
    return [(int(order) >> (3 * p)) & 7 for p in range(n)]
