"""GPU: the k-form glue (N4) drives the block mass kernels: BASELINE config 3's 2-D 1-form element matrix assembled from
the blocks of its component spaces equals the compiled reference's blocks (tests/golden/c3.npz) and is symmetric."""
from pathlib import Path

import numpy as np
import pytest

from conftest import TOL, rel_err
from interplib_b200 import kform as kf
from interplib_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"


def test_one_form_element_matrix_c3(ctx):
    g = dict(np.load(GOLD / "c3.npz"))
    cfg = syn.CONFIGS["c3"]
    basis = [("lagrange-gauss-lobatto", 8)] * 2
    blocks = kf.kform_mass_blocks(ctx, basis, 1, list(cfg.rules), list(cfg.geom_basis), g["geom"], 2)
    assert sorted(blocks) == [(0, 0), (0, 1), (1, 0), (1, 1)]
    for (a, b), m in blocks.items():
        assert rel_err(m, g[f"block_{a}{b}"]) < TOL
    M = kf.assemble_blocks(blocks)
    n = kf.kform_dof_count(basis, 1)
    assert M.shape == (g["geom"].shape[0], n, n) and n == 144
    assert np.abs(M - np.swapaxes(M, 1, 2)).max() <= 1e-13 * np.abs(M).max()  # off-diagonal blocks are separate launches
    # Gram matrix of the union of the two component spaces (scalar products, no metric: SURVEY 8c caveat): positive
    # SEMI-definite -- the spaces share the polynomials of degree <= 7 in both directions
    ev = np.linalg.eigvalsh(M[0])
    assert ev.min() > -1e-12 * ev.max()
