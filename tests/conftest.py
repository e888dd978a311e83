import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """Native artefacts (libll_gpu.so, the oracle port) exist; build them if this is a fresh checkout."""
    import __graft_entry__ as g
    lib = os.path.join(ROOT, "lunar_lander_b200", "libll_gpu.so")
    port = os.path.join(ROOT, "oracle", "_build", "libll_oracle.so")
    host = os.path.join(ROOT, "lunar_lander_b200", "host", "bin", "main_train_d")
    pymods = os.path.join(ROOT, "lunar_lander_b200", "pymodules")
    have_py = os.path.isdir(pymods) and sum(f.endswith(".so") for f in os.listdir(pymods)) >= 2
    if not (os.path.exists(lib) and os.path.exists(port) and os.path.exists(host) and have_py):
        g.build()
    return True


@pytest.fixture(scope="session")
def cfg_path():
    return os.path.join(ROOT, "config.txt")
