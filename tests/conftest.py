"""Shared fixtures.  `-m "not gpu"` runs here (no GPU): oracle vs golden vectors / compiled reference, host
tables, C-ABI symbol export.  `-m gpu` are the parity tests proper: CUDA path (through the C ABI) vs oracle."""
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "ref: needs the compiled reference in oracle/_ref")


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle

    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def refdrive():
    from oracle import build_ref, refdrive as rd

    if not (build_ref.built() or build_ref.available()):
        pytest.skip("compiled reference (oracle/_ref) not available on this machine")
    rd.ref()
    return rd


@pytest.fixture(scope="session")
def ctx():
    import interplib_b200 as ib

    if ib.device_count() == 0:
        pytest.fail("GPU test selected but no CUDA device is visible (the CUDA path has no fallback)")
    c = ib.Context(0)
    yield c
    c.close()


def rel_err(a, ref):
    """SURVEY 8c parity rule: |a-b| <= tol * max(|ref|, ||ref_e||_inf) per element -> returns the worst ratio."""
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert a.shape == ref.shape, (a.shape, ref.shape)
    if a.size == 0:
        return 0.0
    E = a.shape[0]
    diff = np.abs(a - ref).reshape(E, -1)
    scale = np.abs(ref).reshape(E, -1).max(axis=1, keepdims=True)
    scale = np.where(scale > 0, scale, 1.0)
    return float((diff / scale).max())


TOL = 1e-12  # BASELINE.json north_star: relative tolerance for FP64 results
