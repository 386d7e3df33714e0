"""kgwe_b200 — B200-native engine for ONE hot path of asklokesh/k8s-gpu-workload-enhancer:
the scheduler's batched Filter/Prioritize scoring (reference src/scheduler/scheduler.go:114-707).

Contents: csrc/ (CUDA kernels for sm_100a + the C ABI of include/kgwe_engine.h), host/ (C++
mirror of the reference's TopologyAwareScheduler interface over that ABI), and the ctypes
plumbing used by tests and bench.py.  The directory name carries a hyphen, so import it
through `__graft_entry__.load_package()`, which registers it as `kgwe_b200`.
"""
from . import abi, sharded, synth  # noqa: F401
from .engine import Engine, EngineError, unpack_order  # noqa: F401
