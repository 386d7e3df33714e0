"""BASELINE.json configs[1] and configs[4] at their STATED size against the oracle (the reduced-size runs of the
same configs are in test_gpu_parity.py).

configs[1]  1,000 nodes x 1,000 pods, NVLinkOptimal with the minBandwidthGBps mix (reference fields:
            /root/reference/src/scheduler/scheduler.go:206-241, 336-435; types.go:44): 1e6 evaluations - the oracle
            checks every one of them, stateless and stateful.
configs[4]  10,000 nodes, 1,250 gangs x 8 members x 8 GPUs (types.go:415-444; the CRD caps a workload at 64 GPUs,
            deploy/helm/kgwe/crds/gpuworkload-crd.yaml:61-65).  The oracle's sequential queue needs ~25 ms per member
            at this size, so it checks three windows byte for byte - the first gangs, the gangs around the first
            rejected one, the last gangs (each from the ledger the engine's own earlier decisions imply) - and the
            whole queue is checked through its invariants and through the image it leaves behind."""
import subprocess
import sys
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def engine_for(kg, nodes, ledger=None, **kw):
    eng = kg.Engine(max_nodes=len(nodes), **kw)
    eng.upsert_nodes(nodes)
    if ledger is not None:
        eng.seed_ledger(ledger)
    return eng


def test_config2_full_size(kg, orc):
    abi = kg.abi
    nodes, pods, led = kg.synth.config(2)
    assert len(nodes) == 1000 and len(pods) == 1000
    assert (pods["min_bandwidth_gbps"] > 0).sum() > 500 and (pods["arch"] != 0).sum() > 200
    for variant in ("as specified", "minBandwidthGBps = 0 (the pure reference-parity run)"):
        q = pods.copy()
        if variant != "as specified":
            q["min_bandwidth_gbps"] = 0.0
        eng = engine_for(kg, nodes, led)
        for topk in (1, 5, 16):
            s, top = eng.score_batch(q, len(nodes), want_scores=True, topk=topk)
            o, o_top, ctr = orc.score_batch(abi, nodes, led, q, want_scores=True, topk=topk, threads=8)
            assert s.tobytes() == o.tobytes(), variant
            assert top.tobytes() == o_top.tobytes(), variant
        if variant == "as specified":
            assert ((s["flags"] & abi.F_MINBW) != 0).sum() > 1000          # the filter really bites at this size
        st = eng.stats()
        assert st["candidates"] == st["candidates_resolved"] == 3 * ctr["candidates"]
        dec = eng.schedule_queue(q)
        o_dec, o_led, _ = orc.schedule_queue(abi, nodes, led, q)
        assert dec.tobytes() == o_dec.tobytes(), variant
        assert (dec["node_idx"] != abi.NO_NODE).sum() > 100
        # the image after the queue is the oracle's final ledger
        s2, _ = eng.score_batch(q[:32], len(nodes), want_scores=True)
        o2, _, _ = orc.score_batch(abi, nodes, o_led, q[:32], want_scores=True, threads=8)
        assert s2.tobytes() == o2.tobytes(), variant
        eng.close()


def _ledger_after(abi, led, pods, dec, upto):
    """The ledger implied by the initial one and the engine's decisions for pods[:upto]."""
    placed = np.nonzero(dec["node_idx"][:upto] != abi.NO_NODE)[0]
    out = np.zeros(len(led) + len(placed), dtype=abi.ALLOC)
    out[: len(led)] = led
    out["wl_id"][len(led):] = pods["wl_id"][placed]
    out["node_idx"][len(led):] = dec["node_idx"][placed]
    out["gpu_mask"][len(led):] = dec["gpu_mask"][placed]
    return out


def test_config5_full_size_gangs(kg, orc):
    abi = kg.abi
    nodes, pods, led = kg.synth.config(5)
    assert len(nodes) == 10000 and len(pods) == 1250 * 8 and (pods["gpu_count"] == 8).all()
    gang_of = (np.arange(len(pods)) // 8).astype(np.uint32)
    eng = engine_for(kg, nodes, led)
    dec = eng.schedule_gangs(pods, gang_of)
    placed = dec["node_idx"] != abi.NO_NODE
    by_gang = placed.reshape(-1, 8)
    # all-or-nothing; the cluster (80,000 GPUs, some unhealthy or taken) cannot hold all 80,000 requested GPUs
    assert not (by_gang.any(axis=1) & ~by_gang.all(axis=1)).any()
    n_ok = int(by_gang.all(axis=1).sum())
    assert 500 < n_ok < 1250
    first_bad = int(np.argmin(by_gang.all(axis=1)))
    # no GPU handed out twice, nothing on top of the initial ledger, every member on 8 healthy GPUs of one node
    used = {}
    for a in led:
        used[int(a["node_idx"])] = used.get(int(a["node_idx"]), 0) | int(a["gpu_mask"])
    for d in dec[placed]:
        n, m = int(d["node_idx"]), int(d["gpu_mask"])
        assert bin(m).count("1") == 8 and (m & ~int(nodes["healthy_mask"][n])) == 0
        assert used.get(n, 0) & m == 0
        used[n] = used.get(n, 0) | m
    assert eng.stats()["n_allocations"] == len(led) + int(placed.sum())
    # members of rejected gangs that had found a node carry the roll-back flag and nothing else
    rej = ~placed
    assert (dec["gpu_mask"][rej] == 0).all() and ((dec["flags"][rej] & (0xFF ^ abi.F_GANG_ROLLBACK)) == 0).all()
    # three windows byte for byte: the start, around the first rejected gang, the end
    windows = [(0, 40), (max(0, first_bad - 12), min(1250, first_bad + 28)), (1250 - 40, 1250)]
    for g0, g1 in windows:
        p0, p1 = g0 * 8, g1 * 8
        start = _ledger_after(abi, led, pods, dec, p0)
        o_dec, _ = orc.schedule_gangs(abi, nodes, start, pods[p0:p1], gang_of[p0:p1] - gang_of[p0])
        assert dec[p0:p1].tobytes() == o_dec.tobytes(), f"gangs {g0}..{g1}"
    # the image the queue leaves behind (commits of placed gangs in, commits of rejected gangs really gone)
    final = _ledger_after(abi, led, pods, dec, len(pods))
    probe = kg.synth.make_pods(24, seed=99, pref_probs=(0.2, 0.4, 0.0, 0.2, 0.2))
    s, top = eng.score_batch(probe, len(nodes), want_scores=True, topk=3)
    o, o_top, _ = orc.score_batch(abi, nodes, final, probe, want_scores=True, topk=3, threads=8)
    assert s.tobytes() == o.tobytes() and top.tobytes() == o_top.tobytes()
    eng.close()


@pytest.mark.parametrize("path", ["device-side roll-back", "device-side roll-back, grid-wide kernel", "speculative windows"])
def test_gangs_fill_the_cluster(kg, orc, path):
    """The regime configs[4] ends in - most gangs rejected once the cluster is full - on a cluster small enough for
    the oracle to run the whole queue.  Two paths: the queue kernel taking rejected gangs back itself (what
    kgwe_schedule_gangs does whenever the image fits the cluster form or the tile form of the queue kernel - both are run),
    and the host-side speculative windows used for
    larger images, forced here and cut to 64 pods so the queue crosses dozens of window edges, each with rejected
    gangs to undo (a fresh process: the variable is read once)."""
    code = r'''
import sys, numpy as np
sys.path.insert(0, %r)
import __graft_entry__ as entry
kg = entry.load_package()
from oracle import binding as orc
abi = kg.abi
nodes = kg.synth.make_cluster(1000, "mixed", seed=5)
pods = kg.synth.make_pods(250 * 8, seed=5, k_lo=8, k_hi=8, pref_probs=(0, 0.6, 0, 0.2, 0.2))
pods["gpu_count"][::11] = 4            # some gangs with smaller members keep finding room late in the queue
led = kg.synth.make_ledger(nodes, 0.1, seed=5)
gang_of = (np.arange(len(pods)) // 8).astype(np.uint32)
eng = kg.Engine(max_nodes=len(nodes)); eng.upsert_nodes(nodes); eng.seed_ledger(led)
dec = eng.schedule_gangs(pods, gang_of)
o_dec, o_led = orc.schedule_gangs(abi, nodes, led, pods, gang_of)
assert dec.tobytes() == o_dec.tobytes(), "gang decisions differ"
by = (dec["node_idx"] != abi.NO_NODE).reshape(-1, 8)
assert 40 < by.all(axis=1).sum() < 200 and (~by.any(axis=1)).sum() > 50
assert eng.stats()["n_allocations"] == len(o_led)
s, _ = eng.score_batch(pods[:16], len(nodes), want_scores=True)
o, _, _ = orc.score_batch(abi, nodes, o_led, pods[:16], want_scores=True)
assert s.tobytes() == o.tobytes(), "image after the queue differs"
print("OK")
''' % ROOT
    env = dict(os.environ)
    env.pop("KGWE_GANG_WINDOW", None)
    env.pop("KGWE_QUEUE_FORM", None)
    if path.endswith("grid-wide kernel"):
        env["KGWE_QUEUE_FORM"] = "tile"
    if path == "speculative windows":
        env["KGWE_GANG_WINDOW"] = "64"
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


@pytest.mark.parametrize("n_nodes,n_pods", [(16, 300), (1000, 1500), (2100, 1500), (10000, 3000)])
def test_queue_kernel_forms_agree(kg, orc, n_nodes, n_pods):
    """The stateful queue has two device forms: inside one thread-block cluster (kgwe_queue.cuh; 128 lanes per CTA up to
    2,048 nodes, 640 up to 10,240) and grid-wide (schedule_tile_kernel).  Both must give the oracle's decisions, with and
    without hints and minimum bandwidths, and leave the same image and counters behind (KGWE_QUEUE_FORM is read when an
    engine is created)."""
    abi = kg.abi
    nodes = kg.synth.make_cluster(n_nodes, "mixed", seed=21)
    pods = kg.synth.make_pods(n_pods, seed=22)
    pods["min_bandwidth_gbps"][::7] = 300.0
    led = kg.synth.make_ledger(nodes, 0.25, seed=23)
    hint = kg.synth.make_hint(n_pods, n_nodes, 0.01, seed=24)
    got = {}
    old = os.environ.pop("KGWE_QUEUE_FORM", None)
    try:
        for form in ("cluster", "tile"):
            os.environ["KGWE_QUEUE_FORM"] = form
            eng = engine_for(kg, nodes, led)
            d1 = eng.schedule_queue(pods[: n_pods // 2], hint[: n_pods // 2])
            d2 = eng.schedule_queue(pods[n_pods // 2:])
            st = eng.stats()
            probe, _ = eng.score_batch(pods[:8], n_nodes, want_scores=True)
            got[form] = (d1.tobytes(), d2.tobytes(), probe.tobytes(), st["evaluations"], st["candidates_resolved"], st["n_allocations"])
            eng.close()
    finally:
        os.environ.pop("KGWE_QUEUE_FORM", None)
        if old is not None:
            os.environ["KGWE_QUEUE_FORM"] = old
    assert got["cluster"] == got["tile"]
    n_check = n_pods // 2 if n_nodes <= 2100 else 150      # the oracle's sequential queue is slow on 10k nodes
    o_dec = orc.schedule_queue(abi, nodes, led, pods[:n_check], hint[:n_check])[0]
    assert got["cluster"][0][: n_check * abi.PLACEMENT.itemsize] == o_dec.tobytes()
