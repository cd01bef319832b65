"""CPU: incidence matrices and derivative DoFs.  The oracle restatement and the product's host
table layer (ib200_incidence_matrix) are pinned bit-exact against tests/golden/incidence.npz -- outputs of the compiled
reference's incidence_matrix and DegreesOfFreedom.derivative (tests/golden/make_golden_points.py) -- and, where
oracle/_ref is present, against the reference on fresh inputs."""
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
from make_golden_points import ALL_BASIS, DERIV_SPACES  # noqa: E402

GOLD = dict(np.load(Path(__file__).resolve().parent / "golden" / "incidence.npz"))


@pytest.mark.parametrize("bt", ALL_BASIS)
@pytest.mark.parametrize("order", [1, 2, 5, 8, 16])
def test_incidence_matrix_matches_golden(orc, bt, order):
    from interplib_b200 import batch

    ref = GOLD[f"inc_{bt}_{order}"]
    assert np.array_equal(orc.incidence_matrix((bt, order)), ref)
    assert np.array_equal(batch.incidence_matrix((bt, order)), ref)


@pytest.mark.parametrize("name", sorted(DERIV_SPACES))
def test_oracle_dof_derivative_matches_golden(orc, name):
    basis = DERIV_SPACES[name]
    for k in range(len(basis)):
        assert np.array_equal(orc.dof_derivative(basis, GOLD[f"{name}_dofs"], k), GOLD[f"{name}_d{k}"])


def test_oracle_incidence_vs_reference_fresh(orc, refdrive):
    rng = np.random.default_rng(31)
    basis = [("lagrange-uniform", 3), ("legendre", 6), ("bernstein", 1)]
    dofs = rng.standard_normal((3, 4 * 7 * 2))
    for k in range(3):
        assert np.array_equal(orc.dof_derivative(basis, dofs, k), refdrive.dof_derivative(basis, dofs, k))
    for b in [("lagrange-gauss", 7), ("bernstein", 6), ("legendre", 9)]:
        assert np.array_equal(orc.incidence_matrix(b), refdrive.incidence_matrix(b))


def test_incidence_zero_order_is_an_error():
    from interplib_b200 import batch

    with pytest.raises(ValueError):
        batch.incidence_matrix(("legendre", 0))


@pytest.mark.parametrize("bt", ALL_BASIS)
def test_incidence_differentiates_polynomials(orc, bt):
    """tests/test_incidence.py:18-41 without the projection: for a 1-D polynomial in the space, the incidence matrix
    applied to its DoFs reproduces the derivative at arbitrary points (oracle evaluation)."""
    order = 5
    rng = np.random.default_rng(17)
    dofs = rng.standard_normal((1, order + 1))
    x = rng.uniform(-1, 1, (1, 1, 9))
    d_direct = orc.reconstruct_points([(bt, order)], dofs, x, [1])[0, 0]
    d_dofs = orc.incidence_matrix((bt, order)) @ dofs[0]
    d_lower = orc.reconstruct_points([(bt, order - 1)], d_dofs[None], x, [0])[0, 0]
    assert np.abs(d_direct - d_lower).max() < 1e-11 * max(1.0, np.abs(d_direct).max())
