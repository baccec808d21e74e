import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    """The product package (ocean-simulation_b200/), with liboceanfft.so built if it is stale."""
    import oceanpkg
    oceanpkg.build_module().build()
    return oceanpkg.load()


@pytest.fixture(scope="session")
def po():
    """The CPU oracles (checker only)."""
    from oracle import pyoracle
    pyoracle.build(ref=os.path.isdir("/root/reference"))
    return pyoracle


def rel_l2(a, b):
    cplx = np.iscomplexobj(a) or np.iscomplexobj(b)
    a = np.asarray(a, np.complex128 if cplx else np.float64)
    b = np.asarray(b, np.complex128 if cplx else np.float64)
    d = np.linalg.norm(b.ravel())
    return float(np.linalg.norm((a - b).ravel()) / d) if d > 0 else float(np.linalg.norm((a - b).ravel()))


def max_abs_rel(a, b):
    """max |a-b| / max |b| (the 'max-abs <= 1e-4 * max|field|' criterion of SURVEY 8c)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    m = np.abs(b).max()
    return float(np.abs(a - b).max() / m) if m > 0 else float(np.abs(a - b).max())
