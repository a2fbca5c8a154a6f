import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "refslow: re-runs the live reference head to head (build container, PULSE_REFSLOW=1)")


@pytest.fixture(scope="session")
def hr_table():
    from pulselib_b200 import handranks

    return handranks.load_table()


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import pyoracle

    return pyoracle.lib()


@pytest.fixture(scope="session")
def cuda_device():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)
