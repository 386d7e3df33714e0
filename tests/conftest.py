import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def kg():
    return entry.load_package()


@pytest.fixture(scope="session")
def orc():
    entry.build_oracle()
    from oracle import binding
    return binding


@pytest.fixture(scope="session")
def engine_lib(kg):
    entry.build_engine()
    return kg.abi.load_engine_lib()
