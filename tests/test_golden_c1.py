"""Committed regression vectors for BASELINE configs[0] (tests/golden/c1_oracle.npz, made by
tests/golden/make_golden.py from the faithful oracle): the synthetic generator, both oracle
builds and - on the GPU - the engine must all still reproduce them byte for byte."""
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "c1_oracle.npz"))


def _inputs(kg):
    abi = kg.abi
    nodes, pods, led = kg.synth.config(1)
    assert nodes.view(np.uint8).tobytes() == G["nodes"].tobytes() and pods.view(np.uint8).tobytes() == G["pods"].tobytes()
    hint = kg.synth.make_hint(len(pods), len(nodes), 0.1)
    assert hint.tobytes() == G["hint"].tobytes()
    return abi, nodes, pods, led, hint


@pytest.mark.parametrize("mode", [0, 1], ids=["faithful", "indexed"])
def test_oracle_reproduces_golden(kg, orc, mode):
    abi, nodes, pods, led, hint = _inputs(kg)
    s, t, _ = orc.score_batch(abi, nodes, led, pods, hint=hint, want_scores=True, topk=4, mode=mode)
    d, lo, _ = orc.schedule_queue(abi, nodes, led, pods, mode=mode)
    assert s.view(np.uint8).tobytes() == G["scores"].tobytes() and t.view(np.uint8).tobytes() == G["top4"].tobytes()
    assert d.view(np.uint8).tobytes() == G["decisions"].tobytes() and lo.view(np.uint8).tobytes() == G["ledger_out"].tobytes()


@pytest.mark.gpu
def test_engine_reproduces_golden(kg):
    abi, nodes, pods, led, hint = _inputs(kg)
    eng = kg.Engine(max_nodes=len(nodes))
    eng.upsert_nodes(nodes)
    s, t = eng.score_batch(pods, len(nodes), hint=hint, want_scores=True, topk=4)
    assert s.view(np.uint8).tobytes() == G["scores"].tobytes() and t.view(np.uint8).tobytes() == G["top4"].tobytes()
    d = eng.schedule_queue(pods)
    assert d.view(np.uint8).tobytes() == G["decisions"].tobytes()
    eng.close()
