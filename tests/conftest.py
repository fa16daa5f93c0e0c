import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def rel_l2(a, b):
    a = np.asarray(a).ravel()
    b = np.asarray(b).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


@pytest.fixture(scope="session")
def oracle_lib():
    import orc
    orc.build()
    return orc


@pytest.fixture(scope="session", autouse=True)
def _start_fullsize_oracles(request):
    """GPU suite on a GPU box: start the long full-size oracle runs now (thread pool, tests/fullsize.py)
    so that they overlap with the rest of the suite."""
    mark = request.config.getoption("-m") or ""
    if "gpu" in mark and "not gpu" not in mark:
        try:
            import nls_b200
            if nls_b200.load_library().nls_device_count() > 0:
                sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
                import fullsize
                fullsize.start_all()
        except Exception:  # no library / no device: the GPU tests themselves will say so
            pass
    yield
