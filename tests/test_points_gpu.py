"""GPU parity of the arbitrary-point kernels (ib200_evaluate_points_batch /
ib200_reconstruct_points_batch through the C ABI) against the golden vectors of the compiled reference, the CPU oracle
on seeded inputs, and -- at BASELINE config 2's full size -- against K1 at the rule nodes.
Tolerance: |a - b| <= 1e-12 * max(|ref|, ||ref_e||_inf) per element."""
import sys
from pathlib import Path

import numpy as np
import pytest

from conftest import TOL, rel_err
from interplib_b200 import batch
from interplib_b200 import synthetic as syn

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
from make_golden_points import ALL_BASIS, SPACES  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = dict(np.load(Path(__file__).resolve().parent / "golden" / "points.npz"))


def _err(a, ref):
    E = ref.shape[0]
    return rel_err(np.asarray(a).reshape(E, -1), ref.reshape(E, -1))


@pytest.mark.parametrize("name", sorted(SPACES))
def test_points_golden(ctx, name):
    basis = SPACES[name]
    x, dofs = GOLD[f"{name}_x"], GOLD[f"{name}_dofs"]
    assert _err(ctx.evaluate_points(basis, x), GOLD[f"{name}_eval"]) < TOL
    assert _err(ctx.reconstruct_points(basis, dofs, x)[:, 0, :], GOLD[f"{name}_recon"]) < TOL


@pytest.mark.parametrize("bt", ALL_BASIS)
@pytest.mark.parametrize("order", [1, 4, 9])
def test_points_1d_values_and_derivatives_golden(ctx, bt, order):
    x = GOLD["x1"].reshape(1, 1, -1)
    n = order + 1
    val = ctx.evaluate_points([(bt, order)], x)[0]
    assert rel_err(val, GOLD[f"val_{bt}_{order}"]) < TOL
    der = ctx.reconstruct_points([(bt, order)], np.eye(n), np.repeat(x, n, axis=0), [1])[:, 0, :].T
    assert rel_err(der[None], GOLD[f"der_{bt}_{order}"][None]) < TOL


@pytest.mark.parametrize("basis", [
    [("lagrange-gauss-lobatto", 6)] * 2,
    [("lagrange-gauss-lobatto", 16)] * 2,
    [("bernstein", 4)],
    [("legendre", 5), ("lagrange-uniform", 3)],
    [("lagrange-chebyshev-gauss", 3), ("bernstein", 2), ("lagrange-gauss", 2)],
])
@pytest.mark.parametrize("E,P", [(1, 1), (37, 50), (400, 129)])
def test_points_vs_oracle(ctx, orc, basis, E, P):
    rng = np.random.default_rng(E * 1000 + P)
    d = len(basis)
    N = int(np.prod([o + 1 for _, o in basis]))
    x = rng.uniform(-1, 1, (E, d, P))
    dofs = rng.standard_normal((E, N))
    masks = [0] + [1 << k for k in range(d)] + ([3] if d > 1 else [])
    got = ctx.reconstruct_points(basis, dofs, x, masks)
    ref = orc.reconstruct_points(basis, dofs, x, masks)
    for f in range(len(masks)):
        assert _err(got[:, f], ref[:, f]) < TOL, f"field {f}"
    if N * P * E < 5_000_000:
        assert _err(ctx.evaluate_points(basis, x), orc.evaluate_points(basis, x)) < TOL


@pytest.mark.parametrize("orders", [(6, 6), (1, 1), (3, 4), (5, 4), (7, 7), (0, 1)])
@pytest.mark.parametrize("E,P", [(5, 64), (131, 50), (3, 300), (200, 1)])
def test_points_register_kernel_is_bit_identical_to_the_cooperative_one(ctx, orc, orders, E, P, monkeypatch):
    """2-D Lagrange-family spaces with n <= 8 run the thread-per-point register kernel (compile-time shapes); the same
    call with IB200_POINTS_NO_REG goes through the cooperative shared-memory kernel: identical bits, every field mask,
    ragged tiles (E not a multiple of the elements per pass, P not a multiple of the warp) -- and both within TOL of
    the oracle."""
    basis = [("lagrange-gauss-lobatto", orders[0]), ("lagrange-gauss", orders[1])]
    rng = np.random.default_rng(E * 77 + P)
    N = (orders[0] + 1) * (orders[1] + 1)
    x = rng.uniform(-1, 1, (E, 2, P))
    dofs = rng.standard_normal((E, N))
    for masks in ([0], [0, 1, 2], [3, 2, 1, 0, 1], [1], [2, 2]):
        got = ctx.reconstruct_points(basis, dofs, x, masks)
        monkeypatch.setenv("IB200_POINTS_NO_REG", "1")
        coop = ctx.reconstruct_points(basis, dofs, x, masks)
        monkeypatch.delenv("IB200_POINTS_NO_REG")
        assert np.array_equal(got, coop), f"masks {masks}"
        ref = orc.reconstruct_points(basis, dofs, x, masks)
        for f in range(len(masks)):
            assert _err(got[:, f], ref[:, f]) < TOL, f"masks {masks} field {f}"
    # FunctionSpace.evaluate: warp-per-32-points register kernel against the cooperative one and the oracle
    phi = ctx.evaluate_points(basis, x)
    monkeypatch.setenv("IB200_POINTS_NO_REG", "1")
    phi_coop = ctx.evaluate_points(basis, x)
    monkeypatch.delenv("IB200_POINTS_NO_REG")
    assert np.array_equal(phi, phi_coop)
    assert _err(phi, orc.evaluate_points(basis, x)) < TOL


def test_points_empty_and_bad_arguments(ctx):
    basis = [("lagrange-gauss-lobatto", 2)] * 2
    assert ctx.reconstruct_points(basis, np.zeros((0, 9)), np.zeros((0, 2, 5))).shape == (0, 1, 5)
    assert ctx.reconstruct_points(basis, np.zeros((3, 9)), np.zeros((3, 2, 0))).shape == (3, 1, 0)
    assert ctx.evaluate_points(basis, np.zeros((2, 2, 0))).shape == (2, 0, 9)
    from interplib_b200 import InterpError

    with pytest.raises((InterpError, ValueError)):
        ctx.reconstruct_points(basis, np.zeros((1, 9)), np.zeros((1, 2, 4)), [4])  # derivative along a third dimension
    with pytest.raises((InterpError, ValueError)):
        ctx.evaluate_points([("lagrange-uniform", 1)] * 4, np.zeros((1, 4, 2)))  # four dimensions


def test_points_at_rule_nodes_equal_k1_full_size(ctx):
    """2^18 elements of BASELINE config 2: reconstruct_points at the tensor Gauss nodes == K1 (value and gradient)."""
    cfg = syn.CONFIGS["c2"]
    basis, rules = list(cfg.basis), list(cfg.rules)
    E = 1 << 18
    u = np.random.default_rng(3).random((E, cfg.N))
    nodes = batch.integration_space_nodes(rules).reshape(2, -1)  # [d][Q]
    x = np.ascontiguousarray(np.broadcast_to(nodes, (E, 2, cfg.Q)))
    k1 = ctx.reconstruct(basis, rules, u, [0, 1, 2])
    pts = ctx.reconstruct_points(basis, u, x, [0, 1, 2])
    for f in range(3):
        assert _err(pts[:, f], k1[:, f]) < TOL
    # partition of unity at random points, zero gradient
    xr = np.random.default_rng(4).uniform(-1, 1, (1024, 2, 33))
    ones = ctx.reconstruct_points(basis, np.ones((1024, cfg.N)), xr, [0, 1, 2])
    assert np.abs(ones[:, 0] - 1).max() < 1e-13 and np.abs(ones[:, 1:]).max() < 1e-10


def test_points_device_resident(ctx):
    from interplib_b200 import DEVICE

    basis = [("lagrange-gauss-lobatto", 4)] * 2
    rng = np.random.default_rng(8)
    E, P = 5000, 40
    u, x = rng.random((E, 25)), rng.uniform(-1, 1, (E, 2, P))
    ref = ctx.reconstruct_points(basis, u, x, [0, 2])
    d_u, d_x = ctx.alloc(u.nbytes).upload(u), ctx.alloc(x.nbytes).upload(x)
    d_o = ctx.alloc(E * 2 * P * 8)
    ctx.reconstruct_points(basis, d_u.ptr, d_x.ptr, [0, 2], out=d_o.ptr, where=DEVICE, n_elements=E, n_points=P)
    ctx.synchronize()
    assert np.array_equal(d_o.download((E, 2, P)), ref)
    for b in (d_u, d_x, d_o):
        b.free()
