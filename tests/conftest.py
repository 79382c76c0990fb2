import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def native_lib():
    """Builds (if stale) and loads the C-ABI library; never falls back."""
    from trase_b200 import build as tb_build, _lib
    tb_build.build()
    return _lib.load()


GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ["sh_deg3_bg0", "sh_deg1_bg1", "precomp_bg05"]


def load_golden(name):
    import numpy as np
    return np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
