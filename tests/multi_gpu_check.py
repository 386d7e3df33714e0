"""Run under torchrun on N GPUs: node-sharded scoring (NCCL all-gather + merge) must equal the
oracle's ranking over the whole cluster.  `python -m torch.distributed.run --nproc-per-node N
tests/multi_gpu_check.py`.  Exit code 0 = parity on every rank."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    kg = entry.load_package()
    from oracle import binding as orc
    abi, synth = kg.abi, kg.synth
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)

    def check_case(nodes, pods, led, topk, label):
        N, P = len(nodes), len(pods)
        lo, hi = kg.sharded.shard_range(N, rank, world)
        eng = kg.Engine(max_nodes=hi - lo, device=local, node_base=lo)
        eng.upsert_nodes(nodes[lo:hi])
        for a in led:
            if lo <= a["node_idx"] < hi:
                eng.commit(a["wl_id"], a["node_idx"] - lo, a["gpu_mask"])
        sc = kg.sharded.ShardedScorer(eng, hi - lo, world, dist, stream.cuda_stream)
        sc.alloc(torch, dev, P, topk, with_matrix=True)
        d_pods = torch.from_numpy(pods.view(np.uint8).reshape(P, abi.POD_REQ.itemsize)).to(dev)
        merged = sc.step(d_pods)
        torch.cuda.synchronize()
        got = merged.cpu().numpy().view(abi.PLACEMENT).reshape(P, topk)
        local_scores = sc.d_scores.cpu().numpy().view(abi.NODE_SCORE).reshape(P, hi - lo)
        o_scores, want, _ = orc.score_batch(abi, nodes, led, pods, want_scores=True, topk=topk, threads=4)
        ok = got.tobytes() == want.tobytes() and local_scores.tobytes() == np.ascontiguousarray(o_scores[:, lo:hi]).tobytes()
        # the same batch with the exchange fused into the finalisation kernel (P2P stores), top-1, several steps in a row
        sc1 = kg.sharded.ShardedScorer(eng, hi - lo, world, dist, stream.cuda_stream, rank=rank)
        sc1.alloc(torch, dev, P, 1, with_matrix=False)
        p2p_on = sc1.enable_p2p(torch, dev, P)
        _, want1, _ = orc.score_batch(abi, nodes, led, pods, want_scores=False, topk=1, threads=4)
        for form in ("flags", "barrier"):
            # "flags": kgwe_score_merge_p2p (records + flag stores over NVLink, the merge kernel waits on the device);
            # "barrier": finalise + P2P stores, symmetric-memory barrier, merge
            sc1.use_barrier = form == "barrier"
            for it in range(5):
                m = sc1.step(d_pods)
                torch.cuda.synchronize()
                got1 = m.cpu().numpy().view(abi.PLACEMENT).reshape(P, 1)
                ok = ok and got1.tobytes() == want1.tobytes()
            if sc1.flags is not None:
                ok = ok and int(sc1.flags[0][world].item()) == 0          # no merge kernel gave up waiting
        # back-to-back steps without a host synchronisation in between (a rank may run ahead of the others by a step)
        sc1.use_barrier = False
        for it in range(12):
            m = sc1.step(d_pods)
        torch.cuda.synchronize()
        ok = ok and m.cpu().numpy().view(abi.PLACEMENT).reshape(P, 1).tobytes() == want1.tobytes()
        if rank == 0:
            print(f"fused P2P exchange: {'on' if p2p_on else 'unavailable (NCCL path used)'}")
        # stateful queue over the sharded cluster: per-workload all-gather, owner rank commits
        qpods = pods[:120].copy()
        nq = len(qpods)
        qpods["wl_id"][nq - nq // 4:] = qpods["wl_id"][nq // 10: nq // 10 + nq // 4]      # scheduled again: the previous entry may live on another rank

        def release_if_held(wl):
            try:
                eng.release(wl)
            except kg.EngineError as ex:
                assert ex.code == -5          # KGWE_ERR_NOT_FOUND: this rank held nothing for it

        got_q = kg.sharded.schedule_queue_sharded(
            dist, rank, world, (lo, hi), qpods,
            lambda pod, h: eng.score_batch(pod, hi - lo, want_scores=False, topk=1)[1][0, 0],
            lambda wl, n, m: eng.commit(wl, n, m), device=dev, torch=torch, release=release_if_held)
        want_q, _, _ = orc.schedule_queue(abi, nodes, led, qpods)
        ok = ok and got_q.tobytes() == want_q.tobytes()
        eng.close()
        # the same queue exchanged on the device (kgwe_schedule_queue_p2p): fresh engines, two calls (the sequence numbers run
        # on), hints on the first; workload ids must be new, so the repeated tail of qpods is left out
        if p2p_on:
            eng2 = kg.Engine(max_nodes=hi - lo, device=local, node_base=lo)
            eng2.upsert_nodes(nodes[lo:hi])
            for a in led:
                if lo <= a["node_idx"] < hi:
                    eng2.commit(a["wl_id"], a["node_idx"] - lo, a["gpu_mask"])
            sq = kg.sharded.ShardedQueue(eng2, world, rank, dist, torch, dev)
            held = set(int(x) for x in led["wl_id"])
            keep, seen_ids = [], set()
            for i, w in enumerate(pods["wl_id"]):
                if int(w) not in held and int(w) not in seen_ids:
                    keep.append(i); seen_ids.add(int(w))
            fresh = pods[np.array(keep, dtype=np.int64)] if keep else pods[:0]
            half = len(fresh) // 2
            hq = synth.make_hint(half, N, 0.02, seed=77)
            local_hint = np.packbits(np.unpackbits(hq, axis=1, bitorder="little")[:, lo:hi], axis=1, bitorder="little")
            got1 = sq.merged(sq.schedule(fresh[:half], local_hint))
            got2 = sq.merged(sq.schedule(fresh[half:]))
            want12, _, _ = orc.schedule_queue(abi, nodes, led, fresh, np.concatenate([hq, np.zeros((len(fresh) - half, hq.shape[1]), np.uint8)]))
            same = np.concatenate([got1, got2]).tobytes() == want12.tobytes()
            if rank == 0 and not same:
                print(f"multi_gpu_check: device-exchanged queue differs in case {label}")
            ok = ok and same
            eng2.close()
        if rank == 0 and not ok:
            print(f"multi_gpu_check: MISMATCH in case {label}")
        return ok

    nodes, pods, led = synth.config(3, 0.1)
    N, P, topk = len(nodes), 500, 4
    ok = check_case(nodes, pods[:500], led, topk, "config 3 x 0.1")
    # a few cases from the differential fuzzer's input space (tests/fuzz_parity.py), sharded
    import importlib.util
    spec = importlib.util.spec_from_file_location("fuzz_parity", os.path.join(ROOT, "tests", "fuzz_parity.py"))
    fz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fz)
    for seed in (11, 12, 13, 14):
        f_nodes, f_pods, f_led, _, _, f_topk, _, _ = fz.make_case(abi, synth, seed, n_lo=8 * world, n_span=300)
        ok = check_case(f_nodes, f_pods, f_led, max(1, f_topk), f"fuzz seed {seed}") and ok
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"multi_gpu_check world={world}: {'PARITY OK' if flag.item() else 'MISMATCH'} ({P} pods x {N} nodes, top-{topk})")
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
