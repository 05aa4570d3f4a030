import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "gnss-sdr-1_b200"))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def service():
    """One gb200_service for the whole GPU session (fails loudly if the library or the GPU is missing)."""
    import gnss_b200 as gb

    s = gb.Service(device=0, ring_capacity=1 << 23, ring_slack=1 << 21)
    s.tune_from_env()  # GB200_TUNE="key=value,...": lets a child pytest run the same tests under a tuning switch
    yield s
    s.close()
