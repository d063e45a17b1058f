"""pytest configuration: `gpu` marker + shared fixtures.

`-m "not gpu"` runs here (no GPU): oracle vs golden vectors, host logic, ABI checks.
`-m gpu` runs on a B200: parity of the CUDA path (through the C ABI) against the oracle.
"""

import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) device")


@pytest.fixture(scope="session")
def oracle_model():
    from oracle.effnet_b0 import make_calibrated_oracle

    return make_calibrated_oracle(seed=0)


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    return np.load(ROOT / "tests" / "golden" / "effnet_b0_seed0.npz")
