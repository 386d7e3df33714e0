"""Small end-to-end workload for compute-sanitizer runs (memcheck / racecheck): every kernel of the
engine on config 1 (16 nodes) plus a 300-node mixed cluster, checked against the oracle.
  compute-sanitizer --tool memcheck python tests/sanitizer_workload.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

kg = entry.load_package()
from oracle import binding as orc  # noqa: E402

abi, synth = kg.abi, kg.synth
for nodes, pods, led in (synth.config(1), (synth.make_cluster(300, "mixed", seed=3, mig_frac=0.3, partial_node_frac=0.2),
                                           synth.make_pods(96, seed=3, arch_frac=0.3, minbw_choices=(0.0, 500.0)), None)):
    if led is None:
        led = synth.make_ledger(nodes, 0.3, seed=3)
    pods = pods[:96].copy()
    N = len(nodes)
    for flags in (0, abi.CFG_MEMOISE_GROUPS):
        eng = kg.Engine(max_nodes=N, flags=flags)
        eng.upsert_nodes(nodes)
        eng.seed_ledger(led)
        hint = synth.make_hint(len(pods), N, 0.05)
        s, t = eng.score_batch(pods, N, hint=hint, want_scores=True, topk=4)
        o, ot, _ = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=True, topk=4)
        assert s.tobytes() == o.tobytes() and t.tobytes() == ot.tobytes()
        _, t1 = eng.score_batch(pods[:5], N, want_scores=False, topk=1)                 # latency path
        _, o1, _ = orc.score_batch(abi, nodes, led, pods[:5], want_scores=False, topk=1)
        assert t1.tobytes() == o1.tobytes()
        px = pods.copy(); px["flags"] = abi.POD_EXHAUSTIVE
        s, _ = eng.score_batch(px, N, want_scores=True)
        o, _, _ = orc.score_batch(abi, nodes, led, px, want_scores=True)
        assert s.tobytes() == o.tobytes()
        d = eng.schedule_queue(pods)
        od, oled, _ = orc.schedule_queue(abi, nodes, led, pods)
        assert d.tobytes() == od.tobytes()
        assert eng.placement_terms(d).tobytes() == orc.placement_terms(abi, nodes, d).tobytes()
        reqs = synth.make_hint_reqs(40, seed=3)
        assert eng.topology_hints(reqs).tobytes() == orc.topology_hints(abi, nodes, reqs).tobytes()
        assert np.array_equal(eng.topology_gauges(N), orc.topology_gauges(abi, nodes))
        eng.close()
eng = kg.Engine(max_nodes=1)
g, p = synth.make_mig(500, 2000)
a, ga = eng.mig_pack(g, p)
oa, oga = orc.mig_pack(abi, g, p)
assert a.tobytes() == oa.tobytes() and ga.tobytes() == oga.tobytes()
tgt = np.zeros(10); tgt[[1, 3, 5]] = [0.7, 0.2, 0.1]
assert eng.mig_rebalance(g, tgt)[0].tobytes() == orc.mig_rebalance(abi, g, tgt)[0].tobytes()
eng.close()
print("sanitizer workload OK")
