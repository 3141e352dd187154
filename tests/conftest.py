import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def native():
    """Build (if stale) the native libraries in-tree once per session."""
    from spmv_research_b200 import build
    build.build_all()
    return True


@pytest.fixture(scope="session")
def oracle(native):
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_golden.npz"))


@pytest.fixture(scope="session")
def golden_matrices(native):
    """The generator-made matrices of the golden fixture, produced by the product's own generator."""
    import matrices as M
    from spmv_research_b200 import amg
    out = []
    for line in M.GOLDEN_LINES:
        p = amg.MatrixParams.parse(line)
        out.append((p,) + amg.generate(p))
    return out
