"""kgwe_score_batch splits a batch into passes when the score matrix exceeds its scratch chunk
(1 GiB by default).  A subprocess shrinks the chunk to 64 KiB so a small batch needs many passes;
matrix, top-k and hints must still come out identical to the oracle."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = r"""
import sys
sys.path.insert(0, %r)
import numpy as np
import __graft_entry__ as entry
kg = entry.load_package()
from oracle import binding as orc
abi, synth = kg.abi, kg.synth
nodes, pods, led = synth.config(2, 0.12)          # 120 nodes x 16 B = 1920 B per pod row -> 34 pods per 64 KiB pass
pods = pods[:120]
hint = synth.make_hint(len(pods), len(nodes), 0.05)
eng = kg.Engine(max_nodes=len(nodes)); eng.upsert_nodes(nodes); eng.seed_ledger(led)
for topk in (0, 1, 7):
    s, t = eng.score_batch(pods, len(nodes), hint=hint, want_scores=True, topk=topk)
    o, ot, _ = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=True, topk=topk)
    assert s.tobytes() == o.tobytes()
    assert (t is None and ot is None) or t.tobytes() == ot.tobytes()
_, t = eng.score_batch(pods, len(nodes), hint=hint, want_scores=False, topk=5)      # internal matrix scratch, many passes
_, ot, _ = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=False, topk=5)
assert t.tobytes() == ot.tobytes()
print("multipass ok", eng.stats()["kernel_launches"])
""" % ROOT


@pytest.mark.gpu
def test_score_batch_multi_pass():
    env = dict(os.environ, KGWE_SCORE_SCRATCH_BYTES=str(64 * 1024))
    r = subprocess.run([sys.executable, "-c", SCRIPT], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    launches = int(r.stdout.strip().split()[-1])
    assert launches >= 4 * 4        # at least 4 passes per call
