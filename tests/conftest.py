"""Shared fixtures.  `-m "not gpu"` runs the oracle / host-logic tests on CPU; `-m gpu` runs the
parity tests proper, which call the CUDA library through its C-ABI and check it against the
oracle (oracle/ is test infrastructure only)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def port():
    import oracle
    return oracle.Port()


@pytest.fixture(scope="session")
def ref_serial():
    import oracle
    if not oracle.have_ref("serial"):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return oracle.Ref("serial")


@pytest.fixture(scope="session")
def ctx():
    from conjugategradient_b200 import solver
    c = solver.Context(0)  # raises loudly when there is no GPU / no library: no fallback
    yield c
    c.close()


def rel_err(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    d = np.linalg.norm(b)
    return np.linalg.norm(a - b) / (d if d > 0 else 1.0)
