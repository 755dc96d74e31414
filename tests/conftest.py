import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu under gpurun)")


def rel_err(out, ref):
    """max_k ||out_k - ref_k||_inf / max(1, ||ref_k||_inf), k = knot point."""
    import numpy as np

    out = np.asarray(out, dtype=np.float64).reshape(len(ref), -1)
    ref = np.asarray(ref, dtype=np.float64).reshape(len(ref), -1)
    num = np.abs(out - ref).max(axis=1)
    den = np.maximum(1.0, np.abs(ref).max(axis=1))
    return float((num / den).max())


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        return dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))

    return load
