"""pytest configuration: the `gpu` marker and shared helpers.

`-m "not gpu"`: oracle vs golden vectors, host logic, C-ABI symbol check, and the
host simulation of the kernels' per-thread logic (tests/hostsim) — no CUDA.
`-m gpu`: parity tests proper; they call the CUDA path through the C-ABI and
compare with the oracle.  They must not read /root/reference.
"""
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
os.environ.setdefault("OMP_NUM_THREADS", str(os.cpu_count() or 8))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from tests import util
    return util.Oracle()


@pytest.fixture(scope="session")
def hostsim():
    from tests import util
    return util.HostSim()
