"""C++ mirror of the reference's TopologyAwareScheduler interface (host/): packer on CPU, the
full Schedule / ScoreNodes / ReleaseAllocation surface on the GPU against the oracle."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST_TESTS = os.path.join(ROOT, "tests", "host")


def _build(target):
    import __graft_entry__ as entry
    entry.build_engine()
    entry.build_host()
    entry.build_oracle()
    subprocess.check_call(["make", "-s", "-C", HOST_TESTS, target])


def test_packer_and_tables_cpu():
    _build("host_packer_test")
    r = subprocess.run([os.path.join(HOST_TESTS, "host_packer_test")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "host_packer_test OK" in r.stdout


@pytest.mark.gpu
def test_scheduler_mirror_against_oracle_gpu():
    _build("host_scheduler_test")
    r = subprocess.run([os.path.join(HOST_TESTS, "host_scheduler_test")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "host_scheduler_test OK" in r.stdout
