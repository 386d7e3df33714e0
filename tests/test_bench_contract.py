"""bench.py's reference arm (the CPU leg) at a tiny scale: it must run without a GPU and print ONE
JSON line with the contract's keys."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line(orc):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--scale", "0.02", "--steps", "2", "--warmup", "1"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "placement_candidates_scored_per_sec" and d["unit"] == "candidates/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["value"] > 0 and d["vs_baseline"] is None
    assert set(d["cpu_baseline"]) >= {"value", "unit", "cores", "kind", "sample"} and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"]


def test_kgwe_arm_refuses_to_run_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        return
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)


def test_uniform_node_count_of_the_bench_line(kg):
    """bench.py's `link_tiers.headline_cluster.uniform_nodes` (how many records take K1's bit-set walk) on hand-made
    cases: the synthetic cluster is all uniform (also with links down and ragged nodes), a per-link jitter ends that,
    a per-node factor does not, a zero-bandwidth active link does."""
    import importlib.util
    import numpy as np
    spec = importlib.util.spec_from_file_location("kgwe_bench", os.path.join(ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    synth = kg.synth
    nodes = synth.make_cluster(200, "mixed", seed=3, inactive_link_frac=0.1, partial_node_frac=0.3)
    assert b.uniform_nodes(nodes) == 200
    rng = synth.SplitMix64(1)
    per_link = nodes.copy()
    per_link["nvl_bw"] = nodes["nvl_bw"] * (0.5 + rng.uniform(200 * 64).reshape(200, 8, 8))
    has_two_links = (nodes["nvl_active"] != 0).sum(axis=1) >= 2
    assert b.uniform_nodes(per_link) == int((~has_two_links).sum())          # nodes with no active link at all stay "uniform"
    per_node = nodes.copy()
    per_node["nvl_bw"] = nodes["nvl_bw"] * (0.5 + rng.uniform(200))[:, None, None]
    assert b.uniform_nodes(per_node) == 200
    zero = nodes.copy()
    zero["nvl_bw"][7, 0, 1] = 0.0                                            # GPU 0's (active) link to GPU 1 carries nothing
    assert (nodes["nvl_active"][7, 0] >> 1) & 1 and nodes["n_gpus"][7] >= 2
    assert b.uniform_nodes(zero) == 199
