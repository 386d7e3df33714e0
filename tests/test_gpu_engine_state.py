"""Engine state handling on the GPU: the ledger's conflict check, refreshes that change a node's GPU list, the
double-buffered cluster image under concurrent updates, per-thread error messages."""
import threading
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def engine_for(kg, nodes, ledger=None, **kw):
    eng = kg.Engine(max_nodes=len(nodes), **kw)
    eng.upsert_nodes(nodes)
    if ledger is not None:
        eng.seed_ledger(ledger)
    return eng


def test_commit_refuses_gpus_that_are_taken(kg, orc):
    """tryScheduleOnNode re-checks the selected GPUs against every allocation (scheduler.go:634-640); the same workload id
    replaces its own entry (recordAllocation, :699)."""
    abi = kg.abi
    nodes = kg.synth.make_cluster(4, "dgx", seed=3, unhealthy_frac=0.0)
    nodes["n_gpus"][3] = 6
    eng = engine_for(kg, nodes)
    eng.commit(1, 0, 0b00001111)
    with pytest.raises(kg.EngineError) as ei:
        eng.commit(2, 0, 0b00011000)                     # GPU 3 is workload 1's
    assert ei.value.code == -6 and "already allocated" in str(ei.value)
    assert eng.stats()["n_allocations"] == 1
    assert eng.read_node(0)["allocated_mask"] == 0b00001111 and eng.read_node(0)["n_allocs"] == 1
    eng.commit(2, 0, 0b00110000)                         # disjoint: fine
    eng.commit(1, 0, 0b00000011)                         # workload 1 again, shrinking its own set: overwrite, no conflict with itself
    assert eng.read_node(0)["allocated_mask"] == 0b00110011 and eng.read_node(0)["n_allocs"] == 2
    eng.commit(1, 1, 0b11111111)                         # ... and moving to another node
    assert eng.read_node(0)["allocated_mask"] == 0b00110000 and eng.read_node(1)["n_allocs"] == 1
    with pytest.raises(kg.EngineError) as ei:
        eng.commit(9, 3, 0b01000000)                     # node 3 has 6 GPUs
    assert ei.value.code == -1
    # the image agrees with a ledger built from what was accepted
    led = np.zeros(2, dtype=abi.ALLOC)
    led["wl_id"], led["node_idx"], led["gpu_mask"] = [1, 2], [1, 0], [0b11111111, 0b00110000]
    pods = kg.synth.make_pods(12, seed=3)
    s, _ = eng.score_batch(pods, len(nodes), want_scores=True)
    o, _, _ = orc.score_batch(abi, nodes, led, pods, want_scores=True)
    assert s.tobytes() == o.tobytes()
    eng.close()


def _permute_node(abi, node, new_of_old, n_new):
    import importlib.util, os      # the fuzzer's life-cycle mode uses the same helper
    spec = importlib.util.spec_from_file_location("fuzz_parity", os.path.join(os.path.dirname(os.path.abspath(__file__)), "fuzz_parity.py"))
    fz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fz)
    return fz.permute_node(abi, node, new_of_old, n_new)


def test_refresh_that_reorders_gpus_moves_the_allocations(kg, orc):
    """The reference keys allocations by GPU UUID (scheduler.go:586-598); the engine by index, so a refresh that changes a
    node's GPU list carries the old->new index map (kgwe_node_upsert_remap).  Compared with an engine built from scratch
    out of the re-ordered node and the re-indexed ledger, and with the oracle."""
    abi = kg.abi
    nodes = kg.synth.make_cluster(6, "mixed", seed=11, unhealthy_frac=0.1)
    nodes["util"] = np.arange(48).reshape(6, 8) * 2.0            # GPUs are distinguishable
    nodes["free_bytes"] = (np.arange(48).reshape(6, 8) % 7 + 10).astype(np.uint64) << np.uint64(30)
    led = np.zeros(3, dtype=abi.ALLOC)
    led["wl_id"], led["node_idx"], led["gpu_mask"] = [1, 2, 3], [2, 2, 4], [0b00000110, 0b10000000, 0b00001111]
    eng = engine_for(kg, nodes, led)
    # node 2: GPU 1 falls off the bus, the rest is enumerated in reverse order
    new_of_old = [6, -1, 5, 4, 3, 2, 1, 0]
    moved = _permute_node(abi, nodes[2], new_of_old, 7)
    eng.upsert_node_remap(2, moved, new_of_old)
    nodes2 = nodes.copy(); nodes2[2] = moved
    led2 = led.copy()
    led2["gpu_mask"][0] = 1 << 5                     # workload 1 held old GPUs 1 and 2: GPU 1 is gone, GPU 2 is now index 5
    led2["gpu_mask"][1] = 1 << 0                     # old GPU 7 is now index 0
    assert eng.read_node(2)["allocated_mask"] == (1 << 5) | 1 and eng.read_node(2)["n_allocs"] == 2
    pods = kg.synth.make_pods(40, seed=11, pref_probs=(0.2, 0.4, 0.0, 0.2, 0.2))
    s, t = eng.score_batch(pods, len(nodes), want_scores=True, topk=3)
    o, ot, _ = orc.score_batch(abi, nodes2, led2, pods, want_scores=True, topk=3)
    assert s.tobytes() == o.tobytes() and t.tobytes() == ot.tobytes()
    fresh = engine_for(kg, nodes2, led2)
    s2, _ = fresh.score_batch(pods, len(nodes), want_scores=True)
    assert s2.tobytes() == s.tobytes()
    # releasing a moved workload frees the GPU at its NEW index
    eng.release(2)
    assert eng.read_node(2)["allocated_mask"] == 1 << 5
    # a map that is not injective, or that points past the new list, is refused and changes nothing
    for bad in ([0, 0, 1, 2, 3, 4, 5, 6], [7, -1, 5, 4, 3, 2, 1, 0]):
        with pytest.raises(kg.EngineError):
            eng.upsert_node_remap(2, moved, bad)
    assert eng.read_node(2)["allocated_mask"] == 1 << 5
    eng.close(); fresh.close()


def test_image_updates_between_calls_are_seen_exactly(kg, orc):
    """Every scoring call sees exactly the updates made before it (the image is double-buffered on the device: updates go
    to the image that is not being read, then the two swap).  Many small epochs, each checked against the oracle; calls
    alternate between the host path, the latency path and the device path on an external stream."""
    import torch
    abi, synth = kg.abi, kg.synth
    nodes = synth.make_cluster(300, "mixed", seed=8, unhealthy_frac=0.05, inactive_link_frac=0.02)
    pool = synth.make_cluster(300, "mixed", seed=9, unhealthy_frac=0.3, inactive_link_frac=0.2, partial_node_frac=0.3)
    pods = synth.make_pods(24, seed=8, pref_probs=(0.2, 0.4, 0.0, 0.2, 0.2))
    eng = engine_for(kg, nodes)
    rng = synth.SplitMix64(8)
    ledger = {}
    next_wl = [1000]
    present_ok = lambda i: int(nodes["n_gpus"][i]) > 0
    stream = torch.cuda.Stream()
    d_pods = torch.from_numpy(pods.view(np.uint8).reshape(len(pods), abi.POD_REQ.itemsize).copy()).cuda()
    d_scores = torch.empty((len(pods), len(nodes), abi.NODE_SCORE.itemsize), dtype=torch.uint8, device="cuda")
    d_top = torch.empty((len(pods), 1, abi.PLACEMENT.itemsize), dtype=torch.uint8, device="cuda")
    for epoch in range(60):
        for _ in range(int(rng.randint(1, 5, 1)[0])):
            i = int(rng.randint(0, len(nodes) - 1, 1)[0])
            kind = int(rng.randint(0, 2, 1)[0])
            if kind == 0:
                nodes[i] = pool[int(rng.randint(0, len(pool) - 1, 1)[0])]
                eng.upsert_node(i, nodes[i])
                ledger = {w: (n, m & ((1 << int(nodes["n_gpus"][n])) - 1) if n == i else m) for w, (n, m) in ledger.items()}
            elif kind == 1:
                free = ((1 << int(nodes["n_gpus"][i])) - 1) & ~sum(m for (n, m) in ledger.values() if n == i)
                if free and present_ok(i):
                    bit = free & -free
                    next_wl[0] += 1
                    ledger[next_wl[0]] = (i, bit)
                    eng.commit(next_wl[0], i, bit)
            elif ledger:
                w = sorted(ledger)[0]
                eng.release(w); del ledger[w]
        led = np.zeros(len(ledger), dtype=abi.ALLOC)
        for j, (w, (n, m)) in enumerate(sorted(ledger.items())):
            led[j]["wl_id"], led[j]["node_idx"], led[j]["gpu_mask"] = w, n, m
        o, ot, _ = orc.score_batch(abi, nodes, led, pods, want_scores=True, topk=1)
        if epoch % 3 == 0:
            s, t = eng.score_batch(pods, len(nodes), want_scores=True, topk=1)
            assert s.tobytes() == o.tobytes() and t.tobytes() == ot.tobytes(), f"epoch {epoch} (host path)"
        elif epoch % 3 == 1:
            _, t = eng.score_batch(pods[:8], len(nodes), want_scores=False, topk=1)
            assert t.tobytes() == ot[:8].tobytes(), f"epoch {epoch} (latency path)"
        else:
            eng.score_batch_device(d_pods.data_ptr(), len(pods), None, d_scores.data_ptr(), 1, d_top.data_ptr(), stream.cuda_stream)
            stream.synchronize()
            assert d_scores.cpu().numpy().tobytes() == o.tobytes() and d_top.cpu().numpy().tobytes() == ot.tobytes(), f"epoch {epoch} (device path)"
    eng.close()


def test_scoring_is_not_held_up_by_an_update_stream(kg, orc):
    """One thread refreshes nodes about a thousand times a second while another makes per-pod Prioritize calls: results
    stay exact (the refreshes re-send unchanged content, so the expected answer is known) and the latency stays close
    to the quiet figure - applying updates never synchronises the scoring call with a stream."""
    abi, synth = kg.abi, kg.synth
    nodes, pods, led = synth.config(3, 0.2)
    eng = engine_for(kg, nodes, led)
    _, want, _ = orc.score_batch(abi, nodes, led, pods[:64], want_scores=False, topk=1, threads=4)

    def run(calls):
        xs = np.empty(calls)
        for i in range(calls):
            a = time.perf_counter()
            _, t = eng.score_batch(pods[i % 64:i % 64 + 1], len(nodes), want_scores=False, topk=1)
            xs[i] = (time.perf_counter() - a) * 1e6
            assert t.tobytes() == want[i % 64:i % 64 + 1].tobytes()
        return xs
    run(300)
    quiet = run(3000)
    stop = threading.Event()
    sent = [0]

    def updater():
        i = 0
        while not stop.is_set():
            eng.upsert_node(i % len(nodes), nodes[i % len(nodes)])
            sent[0] += 1; i += 1
            time.sleep(0.001)
    th = threading.Thread(target=updater)
    th.start()
    busy = run(3000)
    stop.set(); th.join()
    assert sent[0] > 50
    q99, b99 = np.percentile(quiet, 99), np.percentile(busy, 99)
    print(f"quiet p50/p99 {np.percentile(quiet, 50):.1f}/{q99:.1f} us; with {sent[0]} refreshes p50/p99 {np.percentile(busy, 50):.1f}/{b99:.1f} us")
    assert b99 < 3.0 * q99 + 50.0
    eng.close()


def test_error_messages_are_per_thread(kg):
    eng = kg.Engine(max_nodes=4)
    seen = {}

    def bad(tag, idx):
        for _ in range(200):
            try:
                eng.remove_node(idx)
            except kg.EngineError as ex:
                seen.setdefault(tag, set()).add(str(ex))
    a = threading.Thread(target=bad, args=("a", 1)); b = threading.Thread(target=bad, args=("b", 99))
    a.start(); b.start(); a.join(); b.join()
    assert all("node 1 not present" in m for m in seen["a"]) and all("node 99 out of range" in m for m in seen["b"])
    eng.close()


def test_latency_histogram(kg):
    """kgwe_stats carries the reference's scheduling-latency histogram (prometheus_exporter.go:258-263 bucket edges, in
    microseconds): one observation per kgwe_score_batch call, one per workload of a queue call."""
    abi = kg.abi
    nodes, pods, led = kg.synth.config(1)
    eng = engine_for(kg, nodes, led)
    for i in range(200):
        eng.score_batch(pods[i % 50:i % 50 + 1], len(nodes), want_scores=False, topk=1)
    eng.schedule_queue(pods[:40])
    st = eng.stats()
    assert st["latency_count"] == 240 and sum(st["latency_us_bucket"]) == 240
    assert st["latency_sum_us"] > 0 and st["latency_us_bucket"][-1] < 5                  # next to nothing beyond 5 ms
    assert eng.latency_quantile(0.99) <= 250 and eng.latency_quantile(0.5) <= 100        # the p99 panel's reading of these buckets
    eng.close()
