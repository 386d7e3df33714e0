"""Node-sharded scoring over several GPUs: one process per GPU, rank r owns the
contiguous node range [r*N/R, (r+1)*N/R), scores every pod against its shard, and the ranked
local top-k lists (32-byte placements with GLOBAL node indices) are all-gathered once per batch
and merged identically on every rank (score desc, node index asc).  torch.distributed is the
plumbing (NCCL over NVLink on GPUs; gloo in the CPU tests of the protocol)."""
import numpy as np

from . import abi


def shard_range(n_nodes, rank, world):
    """Contiguous node range of `rank`; keeps the lowest-index tie-break a pure key comparison."""
    return rank * n_nodes // world, (rank + 1) * n_nodes // world


def merge_host(lists, topk):
    """lists: [R, P, topk] placements (numpy) -> [P, topk], via kgwe_topk_merge_host."""
    lists = np.ascontiguousarray(lists, dtype=abi.PLACEMENT)
    R, P = lists.shape[0], lists.shape[1]
    out = np.zeros((P, topk), dtype=abi.PLACEMENT)
    rc = abi.load_engine_lib().kgwe_topk_merge_host(abi.ptr(lists), R, P, topk, abi.ptr(out))
    if rc != 0:
        raise RuntimeError(f"kgwe_topk_merge_host: {rc}")
    return out


class ShardedScorer:
    """Device-resident sharded batch scorer.  `engine` holds this rank's shard (created with
    node_base = shard start).  Buffers are torch uint8 tensors on the engine's GPU."""

    def __init__(self, engine, n_local, world, dist=None, stream_ptr=None, rank=0):
        self.eng, self.n_local, self.world, self.dist, self.sptr, self.rank = engine, n_local, world, dist, stream_ptr, rank
        self.p2p = None            # (symmetric-memory tensor [2, world, P, 32], handle) when the fused exchange is on
        self.step_no = 0
        self._torch, self._ext = None, None      # torch and the engine's stream as a torch stream (set by alloc)
        self.flags = None          # symmetric-memory flag words of the barrier-free exchange
        self.use_barrier = False   # True: the older form (finalise + P2P stores, symmetric-memory barrier, merge)

    def enable_p2p(self, torch, dev, n_pods):
        """Fused exchange for top-1 batches: the finalisation kernel stores straight into every rank's
        gather buffer (torch symmetric memory = peer-mapped allocations), double-buffered, one cross-GPU
        barrier per batch instead of an NCCL all-gather.  Returns False (and stays on NCCL) if
        symmetric memory cannot be set up on this box."""
        try:
            import torch.distributed._symmetric_memory as symm_mem
            buf = symm_mem.empty((2, self.world, n_pods, abi.PLACEMENT.itemsize), dtype=torch.uint8, device=dev)
            hdl = symm_mem.rendezvous(buf, self.dist.group.WORLD)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            if len(ptrs) != self.world:
                return False
            self.p2p = (buf, hdl, ptrs)
            # flags of the barrier-free exchange (kgwe_score_merge_p2p): one word per rank + one "gave up" word
            fl = symm_mem.empty((self.world + 1,), dtype=torch.int32, device=dev)
            fl.zero_()
            fh = symm_mem.rendezvous(fl, self.dist.group.WORLD)
            fptrs = [int(p) for p in fh.buffer_ptrs]
            torch.cuda.synchronize()
            self.dist.barrier()                     # every rank's flags are zero before anyone raises one
            self.flags = (fl, fh, fptrs) if len(fptrs) == self.world else None
            return True
        except Exception:          # noqa: BLE001 - any failure: keep the NCCL path
            self.p2p = None
            return False

    def alloc(self, torch, dev, n_pods, topk, with_matrix=True):
        it = abi.PLACEMENT.itemsize
        self.P, self.topk = n_pods, topk
        self._torch = torch
        self._ext = torch.cuda.ExternalStream(self.sptr, device=dev) if self.sptr else None
        self.d_scores = torch.empty((n_pods, self.n_local, abi.NODE_SCORE.itemsize), dtype=torch.uint8, device=dev) if with_matrix else None
        self.d_top = torch.empty((n_pods, topk, it), dtype=torch.uint8, device=dev)
        self.d_gather = torch.empty((self.world, n_pods, topk, it), dtype=torch.uint8, device=dev) if self.world > 1 else None
        self.d_merged = torch.empty((n_pods, topk, it), dtype=torch.uint8, device=dev) if self.world > 1 else self.d_top

    def step(self, d_pods, d_hint=None):
        """One batch: local scoring (+ matrix), all-gather of local top-k, merge.  Asynchronous.
        The engine's kernels run on the stream given at construction; the cross-GPU barrier and the NCCL all-gather
        run on torch's CURRENT stream, so for the duration of the step that is made the same stream - otherwise
        P2P stores -> barrier -> merge would only be ordered by luck."""
        if self.world > 1 and self._ext is not None:
            with self._torch.cuda.stream(self._ext):
                return self._step(d_pods, d_hint)
        return self._step(d_pods, d_hint)

    def _step(self, d_pods, d_hint=None):
        sc = self.d_scores.data_ptr() if self.d_scores is not None else None
        if self.p2p is not None and self.flags is not None and self.topk == 1 and self.world > 1 and not self.use_barrier:
            buf, hdl, ptrs = self.p2p
            self.step_no += 1
            self.eng.score_merge_p2p(d_pods.data_ptr(), self.P, d_hint.data_ptr() if d_hint is not None else None, sc,
                                     ptrs, self.flags[2], self.rank, self.step_no, self.d_merged.data_ptr(), self.sptr)
            return self.d_merged
        if self.p2p is not None and self.topk == 1 and self.world > 1:
            buf, hdl, ptrs = self.p2p
            slot = self.step_no & 1
            self.step_no += 1
            off = slot * self.world * self.P * abi.PLACEMENT.itemsize
            self.eng.score_batch_device_p2p(d_pods.data_ptr(), self.P, d_hint.data_ptr() if d_hint is not None else None, sc,
                                            [p + off for p in ptrs], self.rank, self.sptr)
            hdl.barrier(channel=slot)          # every rank's stores have landed before anyone merges
            self.eng.topk_merge_device(buf[slot].data_ptr(), self.world, self.P, 1, self.d_merged.data_ptr(), self.sptr)
            return self.d_merged
        self.eng.score_batch_device(d_pods.data_ptr(), self.P, d_hint.data_ptr() if d_hint is not None else None,
                                    sc, self.topk, self.d_top.data_ptr(), self.sptr)
        if self.world > 1:
            self.dist.all_gather_into_tensor(self.d_gather, self.d_top)
            self.eng.topk_merge_device(self.d_gather.data_ptr(), self.world, self.P, self.topk, self.d_merged.data_ptr(), self.sptr)
        return self.d_merged


class ShardedQueue:
    """The stateful queue (Schedule with a commit per workload) over a node-sharded cluster, exchanged on the device:
    one engine per rank holds a shard; every rank calls schedule() with the same workloads; the queue kernels of the ranks
    exchange one 64-bit word per workload through symmetric (peer-mapped) memory and the owner of the winning node commits.
    No collective and no host round trip per workload (schedule_queue_sharded below is the host-driven form)."""
    SEQ_MOD = 0xFFFFFF

    def __init__(self, engine, world, rank, dist, torch, dev):
        import torch.distributed._symmetric_memory as symm_mem
        self.eng, self.world, self.rank, self.dist, self.torch, self.dev = engine, world, rank, dist, torch, dev
        self.recv = symm_mem.empty((world * 4,), dtype=torch.int64, device=dev)
        self.recv.zero_()
        self.hdl = symm_mem.rendezvous(self.recv, dist.group.WORLD)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        if len(self.ptrs) != world:
            raise RuntimeError("symmetric memory: not every rank is reachable")
        torch.cuda.synchronize()
        dist.barrier()                 # every rank's receive words are zero before anyone sends
        self.seq = 0

    def schedule(self, pods, hint=None):
        """Decisions whose node this rank owns (NO_NODE for the others)."""
        out = self.eng.schedule_queue_p2p(pods, self.ptrs, self.rank, self.seq, hint)
        self.seq = (self.seq + len(pods)) % self.SEQ_MOD
        return out

    def merged(self, local):
        """All ranks' decisions put together (one all-gather of the result arrays; for tests and reports)."""
        torch = self.torch
        it = abi.PLACEMENT.itemsize
        mine = torch.from_numpy(local.view(np.uint8).reshape(-1).copy()).to(self.dev)
        allr = torch.empty(self.world * mine.numel(), dtype=torch.uint8, device=self.dev)
        self.dist.all_gather_into_tensor(allr, mine)
        parts = allr.cpu().numpy().view(abi.PLACEMENT).reshape(self.world, len(local))
        out = parts[0].copy()
        for r in range(1, self.world):
            take = parts[r]["node_idx"] != abi.NO_NODE
            assert not (take & (out["node_idx"] != abi.NO_NODE)).any(), "two ranks claim the same workload"
            out[take] = parts[r][take]
        return out


def schedule_queue_sharded(dist, rank, world, shard, pods, score_top1, commit, hint=None, device=None, torch=None, release=None):
    """Stateful Schedule() over a node-sharded cluster: per workload every rank scores its
    shard (top-1, global node index), ONE small all-gather of the 32-byte placements, the identical merge on
    every rank (score desc, node index asc), and the rank that owns the winning node commits it.
      score_top1(pod_array_of_1, hint_row_or_None) -> PLACEMENT record of this shard's best node
      commit(wl_id, local_node_idx, gpu_mask)
      release(wl_id)   (optional) drop this rank's entry for wl_id if it has one; called on EVERY rank before the owner
                       commits, because recordAllocation overwrites by UID (scheduler.go:699) and the workload's
                       previous entry may live in another shard
    `shard` = (lo, hi) node range of this rank.  Exact, but it pays a collective per workload: for clusters
    that fit one GPU (10k nodes = 7.68 MB) kgwe_schedule_queue on one engine is the fast path."""
    lo, hi = shard
    out = np.zeros(len(pods), dtype=abi.PLACEMENT)
    it = abi.PLACEMENT.itemsize
    for i in range(len(pods)):
        local = np.ascontiguousarray(score_top1(pods[i:i + 1], None if hint is None else hint[i:i + 1]), dtype=abi.PLACEMENT).reshape(1)
        if world > 1:
            mine = torch.from_numpy(local.view(np.uint8).reshape(it).copy())
            if device is not None:
                mine = mine.to(device)
            gathered = torch.empty(world * it, dtype=torch.uint8, device=mine.device)
            dist.all_gather_into_tensor(gathered, mine)
            lists = gathered.cpu().numpy().view(abi.PLACEMENT).reshape(world, 1, 1)
        else:
            lists = local.reshape(1, 1, 1)
        best = merge_host(lists, 1)[0, 0]
        out[i] = best
        if best["node_idx"] != abi.NO_NODE:
            if release is not None:
                release(int(pods[i]["wl_id"]))
            if lo <= int(best["node_idx"]) < hi:
                commit(int(pods[i]["wl_id"]), int(best["node_idx"]) - lo, int(best["gpu_mask"]))
    return out
