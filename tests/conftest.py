import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))


@pytest.fixture(scope="session")
def golden():
    return load_golden


def rel_l2(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))


def pytest_sessionfinish(session, exitstatus):
    """decision-flip tallies of the GPU parity tests -> $MMIMO_MARGIN_LOG (tests/parity_util.py)"""
    try:
        import parity_util
        parity_util.dump_stats()
    except Exception:
        pass
    try:
        import ci_util
        ci_util.dump_log()   # z statistics of the BER-vs-reference comparisons -> $MMIMO_CI_LOG
    except Exception:
        pass
