"""CPU: pin the arbitrary-point part of the oracle against tests/golden/points.npz -- outputs of
the compiled reference's FunctionSpace.evaluate, interplib.reconstruct and BasisSpecs.values/.derivatives
(tests/golden/make_golden_points.py) -- and, where oracle/_ref is present, against the reference on fresh inputs."""
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
from make_golden_points import ALL_BASIS, SPACES  # noqa: E402

GOLD = dict(np.load(Path(__file__).resolve().parent / "golden" / "points.npz"))


@pytest.mark.parametrize("name", sorted(SPACES))
def test_oracle_points_match_golden(orc, name):
    basis = SPACES[name]
    x, dofs = GOLD[f"{name}_x"], GOLD[f"{name}_dofs"]
    # the tensor basis is a chain of single products: bit-equal to the reference
    assert np.array_equal(orc.evaluate_points(basis, x), GOLD[f"{name}_eval"])
    # reconstruct: NumPy adds the N terms pairwise, the oracle in flat order
    rec = orc.reconstruct_points(basis, dofs, x, [0])[:, 0, :]
    ref = GOLD[f"{name}_recon"]
    assert np.abs(rec - ref).max() <= 1e-14 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("bt", ALL_BASIS)
@pytest.mark.parametrize("order", [1, 4, 9])
def test_oracle_1d_values_and_derivatives_match_golden(orc, bt, order):
    x1 = GOLD["x1"]
    x = x1.reshape(1, 1, -1)
    val = orc.evaluate_points([(bt, order)], x)[0]
    assert np.array_equal(val, GOLD[f"val_{bt}_{order}"])
    # derivative of basis i = reconstruct of the unit DoF vector e_i with the derivative mask (0 + d*1 is exact)
    n = order + 1
    eye = np.eye(n)
    der = np.stack([orc.reconstruct_points([(bt, order)], eye[i:i + 1], x, [1])[0, 0] for i in range(n)], axis=1)
    assert np.array_equal(der, GOLD[f"der_{bt}_{order}"])


def test_oracle_points_vs_reference_fresh(orc, refdrive):
    rng = np.random.default_rng(99)
    basis = [("lagrange-gauss-lobatto", 5), ("bernstein", 2)]
    x = rng.uniform(-1, 1, (4, 2, 11))
    dofs = rng.standard_normal((4, 18))
    assert np.array_equal(orc.evaluate_points(basis, x), refdrive.evaluate_points(basis, x))
    a, b = orc.reconstruct_points(basis, dofs, x)[:, 0], refdrive.reconstruct_points(basis, dofs, x)
    assert np.abs(a - b).max() <= 1e-14 * np.abs(b).max()


def test_host_tables_agree_with_oracle_points(orc):
    """The product's host layer (ib200_function_space_evaluate, A13) against the oracle restatement."""
    from interplib_b200 import batch

    rng = np.random.default_rng(5)
    basis = [("legendre", 3), ("lagrange-uniform", 4)]
    x = rng.uniform(-1, 1, (1, 2, 13))
    host = batch.function_space_evaluate(basis, [x[0, 0], x[0, 1]])
    assert np.array_equal(np.asarray(host).reshape(13, -1), orc.evaluate_points(basis, x)[0])
