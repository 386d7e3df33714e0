"""Regenerates tests/golden/c1_oracle.npz: outputs of the CPU oracle (faithful build) on BASELINE
configs[0] (16-node DGX cluster, 100 pods).  The reference itself cannot run here (Go), so these
are regression vectors for the restatement, not outputs of the reference; the hand-derived
known answers are in kats.json.  Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

kg = entry.load_package()
entry.build_oracle()
from oracle import binding as orc  # noqa: E402

nodes, pods, led = kg.synth.config(1)
hint = kg.synth.make_hint(len(pods), len(nodes), 0.1)
scores, top, _ = orc.score_batch(kg.abi, nodes, led, pods, hint=hint, want_scores=True, topk=4, mode=orc.FAITHFUL)
dec, led_out, _ = orc.schedule_queue(kg.abi, nodes, led, pods, mode=orc.FAITHFUL)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "c1_oracle.npz"),
                    nodes=nodes.view(np.uint8), pods=pods.view(np.uint8), hint=hint,
                    scores=scores.view(np.uint8), top4=top.view(np.uint8), decisions=dec.view(np.uint8), ledger_out=led_out.view(np.uint8))
print("wrote c1_oracle.npz", scores.shape, top.shape, dec.shape)
