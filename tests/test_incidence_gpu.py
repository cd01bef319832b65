"""GPU parity of the derivative-DoF kernel (ib200_dof_derivative_batch through the C ABI):
bit-exact against the golden vectors of the compiled reference and against the oracle, plus the reference's own test
properties (tests/test_incidence.py) in batched form."""
import sys
from pathlib import Path

import numpy as np
import pytest

from interplib_b200 import batch

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
from make_golden_points import ALL_BASIS, DERIV_SPACES  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = dict(np.load(Path(__file__).resolve().parent / "golden" / "incidence.npz"))


@pytest.mark.parametrize("name", sorted(DERIV_SPACES))
def test_dof_derivative_golden_bit_exact(ctx, name):
    basis = DERIV_SPACES[name]
    for k in range(len(basis)):
        assert np.array_equal(ctx.dof_derivative(basis, GOLD[f"{name}_dofs"], k), GOLD[f"{name}_d{k}"])


@pytest.mark.parametrize("basis", [
    [("lagrange-gauss-lobatto", 16)] * 2,
    [("bernstein", 7)],
    [("legendre", 6), ("legendre", 5)],
    [("lagrange-uniform", 2), ("bernstein", 3), ("lagrange-gauss", 4)],
])
@pytest.mark.parametrize("E", [1, 257, 5000])
def test_dof_derivative_vs_oracle_bit_exact(ctx, orc, basis, E):
    N = int(np.prod([o + 1 for _, o in basis]))
    dofs = np.random.default_rng(E).standard_normal((E, N))
    for k in range(len(basis)):
        assert np.array_equal(ctx.dof_derivative(basis, dofs, k), orc.dof_derivative(basis, dofs, k))


@pytest.mark.parametrize("order", [1, 2, 6, 11, 12])
@pytest.mark.parametrize("E", [1, 2, 63, 4097])
def test_dof_derivative_kernel_variants_bit_identical(ctx, orc, order, E, monkeypatch):
    """Orders up to 12 run the line-per-thread kernel with bulk-copied tiles (full tiles through cp.async.bulk, the
    ragged tail through plain loads); IB200_DERIV_NO_COLS sends the same call to the tiled kernel.  Same bits, and the
    oracle's, for every family and both dimensions."""
    for bt in ("lagrange-gauss-lobatto", "legendre", "bernstein"):
        basis = [(bt, order), ("lagrange-gauss", 3)]
        N = (order + 1) * 4
        dofs = np.random.default_rng(E + order).standard_normal((E, N))
        for k in range(2):
            got = ctx.dof_derivative(basis, dofs, k)
            monkeypatch.setenv("IB200_DERIV_NO_COLS", "1")
            tiled = ctx.dof_derivative(basis, dofs, k)
            monkeypatch.delenv("IB200_DERIV_NO_COLS")
            assert np.array_equal(got, tiled), (bt, k)
            assert np.array_equal(got, orc.dof_derivative(basis, dofs, k)), (bt, k)


@pytest.mark.parametrize("bt", ALL_BASIS)
def test_derivative_is_incidence_matrix_times_dofs(ctx, bt):
    """tests/test_incidence.py:44-60: DegreesOfFreedom.derivative == incidence @ values (1-D), here for 2^16 elements."""
    order = 6
    E = 1 << 16
    dofs = np.random.default_rng(3).standard_normal((E, order + 1))
    inc = batch.incidence_matrix((bt, order))
    got = ctx.dof_derivative([(bt, order)], dofs, 0)
    ref = dofs @ inc.T
    assert np.abs(got - ref).max() <= 1e-13 * np.abs(ref).max()


def test_derivative_commutes_with_point_evaluation_c2_full_size(ctx):
    """2^20 elements of BASELINE config 2's space: evaluating the derivative DoFs (order 5 along the dimension) at random
    points equals the derivative-mask reconstruction of the original DoFs (N2 x N3 consistency)."""
    basis = [("lagrange-gauss-lobatto", 6)] * 2
    E = 1 << 20
    rng = np.random.default_rng(9)
    u = rng.random((E, 49))
    x = rng.uniform(-1, 1, (E, 2, 2))
    for k in range(2):
        lower = [(t, o - 1 if d == k else o) for d, (t, o) in enumerate(basis)]
        du = ctx.dof_derivative(basis, u, k)
        a = ctx.reconstruct_points(lower, du, x, [0])
        b = ctx.reconstruct_points(basis, u, x, [1 << k])
        assert np.abs(a - b).max() <= 1e-11 * np.abs(b).max()


def test_dof_derivative_errors_and_empty(ctx):
    basis = [("legendre", 0), ("bernstein", 2)]
    with pytest.raises(ValueError):
        ctx.dof_derivative(basis, np.zeros((1, 3)), 0)  # order 0 along the dimension
    with pytest.raises(ValueError):
        ctx.dof_derivative(basis, np.zeros((1, 3)), 2)  # dimension out of range
    assert ctx.dof_derivative(basis, np.zeros((0, 3)), 1).shape == (0, 2)


def test_interp_mirror_derivative_and_incidence():
    """Drop-in names: DegreesOfFreedom.derivative(idim) and incidence_matrix(BasisSpecs) (tests/test_incidence.py:44-60)."""
    from interplib_b200.interp import BasisSpecs, DegreesOfFreedom, FunctionSpace, incidence_matrix

    specs = BasisSpecs("lagrange-gauss-lobatto", 5)
    fs = FunctionSpace(specs, BasisSpecs("bernstein", 2))
    vals = np.random.default_rng(1).standard_normal((6, 3))
    dof = DegreesOfFreedom(fs, vals)
    d0 = dof.derivative(0)
    assert d0.function_space == fs.lower_order(0) and d0.values.shape == (5, 3)
    assert np.allclose(d0.values, incidence_matrix(specs) @ vals, rtol=1e-13, atol=1e-13)
    with pytest.raises(ValueError):
        dof.derivative(2)
    with pytest.raises(ValueError):
        incidence_matrix(BasisSpecs("legendre", 0))
