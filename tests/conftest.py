"""Test configuration: `-m "not gpu"` runs on the CPU-only build box, `-m gpu` on a B200."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "population-sv-simulator_b200")
for p in (ROOT, PKG, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def oracle_lib():
    """ctypes handle on oracle/_build/liboracle.so (CPU restatement; test infrastructure only)."""
    import oracle_helpers
    return oracle_helpers.load_oracle()


@pytest.fixture(scope="session")
def oracle_bin():
    import oracle_helpers
    oracle_helpers.ensure_oracle_built()
    return os.path.join(ROOT, "oracle", "_build", "synthpop_oracle")


@pytest.fixture(scope="session")
def sp():
    """The product library through its ctypes mirror (fails loudly if the CUDA extension is missing)."""
    from synthpop import lib
    lib.load()
    return lib
