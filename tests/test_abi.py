"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol the
header declares, layouts match, and the product path fails loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_every_declared_symbol(kg, engine_lib):
    header = open(os.path.join(ROOT, "include", "kgwe_engine.h")).read()
    declared = set(re.findall(r"\b(kgwe_[a-z0-9_]+)\s*\(", header))
    assert declared == set(kg.abi.ENGINE_SYMBOLS), declared ^ set(kg.abi.ENGINE_SYMBOLS)
    for name in declared:
        assert hasattr(engine_lib, name), name
    assert b"sm_100a" in engine_lib.kgwe_version()


def test_struct_layouts(kg):
    abi = kg.abi
    assert abi.NODE_IMAGE.itemsize == 768 and abi.NODE_IMAGE.fields["n_allocs"][1] == 760
    assert abi.POD_REQ.itemsize == 32 and abi.NODE_SCORE.itemsize == 16 and abi.PLACEMENT.itemsize == 32
    assert C.sizeof(abi.Config) == 48 and C.sizeof(abi.Stats) == 200
    cfg = abi.Config()
    kg.abi.load_engine_lib().kgwe_config_default(C.byref(cfg))
    assert (cfg.struct_size, cfg.topology_weight, cfg.resource_weight, cfg.balance_weight, cfg.hint_bonus) == (48, 40, 35, 25, 10)
    assert cfg.nvlink_norm_gbps == 900.0


def test_no_cpu_fallback(kg, engine_lib):
    """Without a CUDA device the engine refuses to exist; nothing routes to the oracle."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(kg.EngineError) as ei:
        kg.Engine(max_nodes=16)
    assert ei.value.code == -3
    src = "".join(open(os.path.join(ROOT, "k8s-gpu-workload-enhancer_b200", f)).read() for f in ("abi.py", "engine.py", "__init__.py"))
    assert "oracle" not in src.replace("no CPU fallback", "")


def test_invalid_config_rejected(kg, engine_lib):
    h = C.c_void_p()
    cfg = kg.abi.default_config(0)
    assert engine_lib.kgwe_engine_create(C.byref(cfg), C.byref(h)) == -1
    cfg = kg.abi.default_config(16)
    cfg.nvlink_norm_gbps = 1.0
    assert engine_lib.kgwe_engine_create(C.byref(cfg), C.byref(h)) == -1


def test_synth_is_deterministic(kg):
    a = kg.synth.config(2, 0.05)
    b = kg.synth.config(2, 0.05)
    for x, y in zip(a, b):
        assert x.tobytes() == y.tobytes()
    nodes = a[0]
    assert set(np.unique(nodes["arch"])) <= {kg.abi.ARCH_HOPPER, kg.abi.ARCH_AMPERE, kg.abi.ARCH_BLACKWELL}
