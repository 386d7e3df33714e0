"""The N>1 protocol on CPU: 2 processes over gloo.  Each rank ranks its node shard (the oracle
stands in for the GPU engine here, there is no CUDA on this box), the ranked lists carry global
node indices, one all_gather, and every rank merges with kgwe_topk_merge_host.  The merged list
must equal the oracle's ranking over the whole cluster, bit for bit."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import __graft_entry__ as entry
    kg = entry.load_package()
    from oracle import binding as orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(2, 0.15)
    pods = pods[:64]
    topk = 5
    lo, hi = kg.sharded.shard_range(len(nodes), rank, world)
    local_led = led[(led["node_idx"] >= lo) & (led["node_idx"] < hi)].copy()
    local_led["node_idx"] -= lo
    cfg = abi.default_config(hi - lo, node_base=lo)
    _, top, _ = orc.score_batch(abi, nodes[lo:hi], local_led, pods, want_scores=False, topk=topk, cfg=cfg)
    mine = torch.from_numpy(top.view(np.uint8).reshape(len(pods), topk, abi.PLACEMENT.itemsize).copy())
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    lists = np.stack([g.numpy() for g in gathered]).view(abi.PLACEMENT).reshape(world, len(pods), topk)
    merged = kg.sharded.merge_host(lists, topk)
    _, want, _ = orc.score_batch(abi, nodes, led, pods, want_scores=False, topk=topk)
    ok = merged.tobytes() == want.tobytes()
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok, int((want["node_idx"] >= hi - lo).sum()) if rank == 0 else 0))


def _queue_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import __graft_entry__ as entry
    kg = entry.load_package()
    from oracle import binding as orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(1)                     # 16 nodes: pods compete for the same nodes across shards
    pods = pods[:60].copy()
    pods["wl_id"][40:] = pods["wl_id"][5:25]               # workloads scheduled again: the old entry may sit in the OTHER shard
    lo, hi = kg.sharded.shard_range(len(nodes), rank, world)
    state = {"led": [tuple(x) for x in led[(led["node_idx"] >= lo) & (led["node_idx"] < hi)][["wl_id", "node_idx", "gpu_mask"]].tolist()]}
    cfg = abi.default_config(hi - lo, node_base=lo)

    def score_top1(pod, hint_row):                         # the oracle stands in for this rank's GPU engine
        ledger = np.zeros(len(state["led"]), dtype=abi.ALLOC)
        for i, (w, n, m) in enumerate(state["led"]):
            ledger[i]["wl_id"], ledger[i]["node_idx"], ledger[i]["gpu_mask"] = w, n, m
        _, top, _ = orc.score_batch(abi, nodes[lo:hi], ledger, pod, want_scores=False, topk=1, cfg=cfg)
        return top[0, 0]

    def commit(wl, local_node, mask):
        state["led"].append((wl, local_node, mask))

    def release(wl):
        state["led"] = [x for x in state["led"] if x[0] != wl]

    got = kg.sharded.schedule_queue_sharded(dist, rank, world, (lo, hi), pods, score_top1, commit, torch=torch, release=release)
    want, _, _ = orc.schedule_queue(abi, nodes, led, pods)
    ok = got.tobytes() == want.tobytes()
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok, len(state["led"])))


def test_two_rank_sharded_stateful_queue_matches_single_scheduler(kg, engine_lib):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + os.getpid() % 1000
    procs = [ctx.Process(target=_queue_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
    assert all(n > 0 for _, _, n in res)                   # both shards received commits


def test_two_rank_sharded_topk_merge_matches_global_ranking(kg, engine_lib):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 1000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
    assert max(n for _, _, n in res) > 0          # winners really come from both shards


def test_shard_ranges_cover_all_nodes(kg):
    for n in (1, 7, 16, 1000, 10000):
        for world in (1, 2, 3, 4, 8):
            r = [kg.sharded.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))


def test_merge_host_edge_cases(kg, engine_lib):
    abi = kg.abi
    lists = np.zeros((3, 2, 4), dtype=abi.PLACEMENT)
    lists["node_idx"] = abi.NO_NODE
    # pod 0: shard 0 has (50, n3), (40, n9); shard 2 has (50, n1); pod 1: nothing anywhere
    lists[0, 0, 0]["node_idx"], lists[0, 0, 0]["total"] = 3, 50
    lists[0, 0, 1]["node_idx"], lists[0, 0, 1]["total"] = 9, 40
    lists[2, 0, 0]["node_idx"], lists[2, 0, 0]["total"] = 1, 50
    out = kg.sharded.merge_host(lists, 4)
    assert out[0]["node_idx"].tolist() == [1, 3, 9, abi.NO_NODE]      # tie on 50 -> lower node index first
    assert (out[1]["node_idx"] == abi.NO_NODE).all()
