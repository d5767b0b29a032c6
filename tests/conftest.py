import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def unity_data():
    from bear_b200 import data_file
    return data_file.load(GOLDEN / "unity_brirs_decorrelators.tf")


@pytest.fixture(scope="session")
def small_data():
    """Synthetic data file of the default layout with fewer views and shorter BRIRs."""
    from bear_b200 import data_file
    return data_file.from_tree(data_file.synthetic_tree(seed=1, num_views=12, brir_length=700))


@pytest.fixture(scope="session")
def oracle():
    from oracle import bear_ref
    bear_ref.lib("f32")
    bear_ref.lib("f64")
    return bear_ref


def rel_err_db(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return 10 * np.log10(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-300) + 1e-300)
