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


def test_one_form_element_matrix_single_launch_c3(ctx):
    """ib200_kform_mass_mapped_batch: the whole [144][144] matrix of config 3 from ONE launch (Kronecker kernel) equals the
    compiled reference's blocks, is EXACTLY symmetric (M_10 and M_01^T are the same operations on the same operands), and
    equals the block-by-block path (sum-factorised kernels + strided copies) that serves non-bilinear elements."""
    g = dict(np.load(GOLD / "c3.npz"))
    cfg = syn.CONFIGS["c3"]
    basis = [("lagrange-gauss-lobatto", 8)] * 2
    rules, gb = list(cfg.rules), list(cfg.geom_basis)
    l0 = ctx.stats()["kernel_launches"]
    M = kf.kform_mass_matrix(ctx, basis, 1, rules, gb, g["geom"], 2)
    assert ctx.stats()["kernel_launches"] - l0 == 1
    assert M.shape == (g["geom"].shape[0], 144, 144)
    for a in range(2):
        for b in range(2):
            assert rel_err(M[:, 72 * a:72 * a + 72, 72 * b:72 * b + 72], g[f"block_{a}{b}"]) < TOL
    assert np.array_equal(M, np.swapaxes(M, 1, 2))
    ctx.set_kron(False)
    try:
        l0 = ctx.stats()["kernel_launches"]
        Mb = kf.kform_mass_matrix(ctx, basis, 1, rules, gb, g["geom"], 2)
        assert ctx.stats()["kernel_launches"] - l0 == 4
    finally:
        ctx.set_kron(True)
    assert rel_err(M, Mb) < TOL


@pytest.mark.parametrize("k", [0, 1, 2])
@pytest.mark.parametrize("basis", [[("lagrange-gauss-lobatto", 3), ("bernstein", 5)], [("legendre", 4), ("lagrange-uniform", 2)]])
def test_kform_matrix_equals_blocks(ctx, basis, k):
    """Every 2-D k-form, mixed families and orders, ragged batch: the single-call matrix against the per-block API."""
    rng = np.random.default_rng(11 + k)
    E = 133
    rules, gb = [("gauss", 6), ("gauss", 7)], [("lagrange-uniform", 1)] * 2
    geom = syn.quad_mesh_geometry(144, seed=9).reshape(144, 2, 4)[:E] + 1e-3 * rng.standard_normal((E, 2, 4))
    M = kf.kform_mass_matrix(ctx, basis, k, rules, gb, geom, 2)
    ref = kf.assemble_blocks(kf.kform_mass_blocks(ctx, basis, k, rules, gb, geom, 2))
    assert M.shape == ref.shape and rel_err(M, ref) < TOL
    ctx.set_kron(False)
    try:
        assert rel_err(kf.kform_mass_matrix(ctx, basis, k, rules, gb, geom, 2), ref) < TOL
    finally:
        ctx.set_kron(True)
