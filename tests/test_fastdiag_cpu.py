"""CPU: the host tables of the fast-diagonalisation projection (csrc/fastdiag.c) and the identity the GPU kernel relies
on -- on a bilinear quad M^-1 = (V0 x V1) diag(1 / (alpha + beta lambda0 + gamma lambda1)) (V0 x V1)^T -- checked against
the oracle's mass matrices and right-hand sides (restatement of mass_matrices.c:161-339 / integration.py:106-223)."""
import numpy as np
import pytest

from interplib_b200 import batch
from interplib_b200 import synthetic as syn

GLL, GAUSS = "lagrange-gauss-lobatto", "gauss"


@pytest.mark.parametrize("basis,rule", [
    ((GLL, 4), (GAUSS, 5)), ((GLL, 1), (GAUSS, 1)), (("legendre", 3), (GAUSS, 4)), (("bernstein", 5), ("gauss-lobatto", 7)),
    ((GLL, 16), (GAUSS, 17)), (("lagrange-uniform", 8), (GAUSS, 9)), (("lagrange-gauss", 6), (GAUSS, 6)),
    (("lagrange-chebyshev-gauss", 5), (GAUSS, 8)),
])
def test_tables_diagonalise_the_pencil(basis, rule):
    V, lam, Vw = batch.fastdiag_tables(basis, rule)
    x, w = batch.rule_1d(*rule)[:2]
    B = batch.basis_table_1d(basis, rule)[0]
    A, X = (B * w) @ B.T, (B * w * x) @ B.T
    n = basis[1] + 1
    assert np.abs(V.T @ A @ V - np.eye(n)).max() < 5e-15 * np.linalg.cond(A) ** 0.5
    assert np.abs(V.T @ X @ V - np.diag(lam)).max() < 5e-15 * np.linalg.cond(A) ** 0.5
    assert np.abs(Vw - (V.T @ B) * w).max() < 1e-15 * max(1.0, np.abs(Vw).max())
    # the eigenvalues are points of the reference interval (for IntegrationSpecs(1) its end points, the rule's nodes):
    # the scaling alpha + beta lambda0 + gamma lambda1 is det J at a point of the element, positive on a valid quad
    assert np.all(np.abs(lam) <= 1.0 + 1e-15)


@pytest.mark.parametrize("basis", ["legendre", "bernstein", "lagrange-gauss", GLL, "lagrange-uniform"])
@pytest.mark.parametrize("rule", [GAUSS, "gauss-lobatto"])
@pytest.mark.parametrize("n_basis,n_nodes", [(6, 4), (4, 3), (9, 8), (17, 12)])
def test_under_integrated_space_is_refused(basis, rule, n_basis, n_nodes):
    """Fewer rule nodes than basis functions -> singular 1-D mass matrix: always INTERP_ERROR_BAD_SYSTEM, never tables built
    on a roundoff pivot.  (As many nodes as functions is accepted: test_tables_diagonalise_the_pencil has such a pair.)"""
    with pytest.raises(Exception, match="BAD_SYSTEM"):
        batch.fastdiag_tables((basis, n_basis - 1), (rule, n_nodes - 1))


def fastdiag_project(basis, rules, geom_basis, geom, f):
    """NumPy statement of kernels_fastdiag.cu (same tables, same contractions)."""
    (V0, l0, Vw0), (V1, l1, Vw1) = (batch.fastdiag_tables(b, r) for b, r in zip(basis, rules))
    x0, x1 = (batch.rule_1d(*r)[0] for r in rules)
    lin = []
    for gb in geom_basis:
        v = batch.basis_at_points(gb, np.array([0.0, 1.0]))  # [cnt][n]
        lin.append((v[0], v[1] - v[0]))
    (a0, s0), (a1, s1) = lin
    G = np.stack([np.outer(s0, a1).ravel(), np.outer(a0, s1).ravel(), np.outer(s0, s1).ravel()])
    c = geom.reshape(len(geom), 2, 4)
    d = np.einsum("kc,epc->epk", G, c)  # [E][x|y][xi0, xi1, xi0 xi1]
    alpha = d[:, 0, 0] * d[:, 1, 1] - d[:, 0, 1] * d[:, 1, 0]
    beta = d[:, 0, 0] * d[:, 1, 2] - d[:, 0, 2] * d[:, 1, 0]
    gamma = d[:, 0, 2] * d[:, 1, 1] - d[:, 0, 1] * d[:, 1, 2]
    E = len(geom)
    m0, m1 = len(x0), len(x1)
    det = alpha[:, None, None] + beta[:, None, None] * x0[None, :, None] + gamma[:, None, None] * x1[None, None, :]
    bt = np.einsum("iq,jr,eqr->eij", Vw0, Vw1, f.reshape(E, m0, m1) * det)
    D = alpha[:, None, None] + beta[:, None, None] * l0[None, :, None] + gamma[:, None, None] * l1[None, None, :]
    return np.einsum("ai,bj,eij->eab", V0, V1, bt / D).reshape(E, -1), det.reshape(E, -1)


@pytest.mark.parametrize("basis,rules,gb", [
    ([(GLL, 4)] * 2, [(GAUSS, 5)] * 2, [("lagrange-uniform", 1)] * 2),                       # BASELINE config 5
    ([(GLL, 6)] * 2, [(GAUSS, 7)] * 2, [("bernstein", 1)] * 2),
    ([("legendre", 3)] * 2, [(GAUSS, 4)] * 2, [("legendre", 1), ("lagrange-gauss", 1)]),     # other linear geometry bases
    ([(GLL, 16)] * 2, [(GAUSS, 17)] * 2, [("lagrange-uniform", 1)] * 2),                      # BASELINE config 4
])
def test_identity_against_the_oracle(orc, basis, rules, gb):
    rng = np.random.default_rng(5)
    E = 3 if basis[0][1] > 8 else 24
    # corner coordinates as DoFs of the geometry basis: sample a jittered bilinear map at the basis' own nodes is not
    # needed -- any DoF values define a bilinear map
    geom = syn.quad_mesh_geometry(64, seed=3).reshape(64, 2, 2, 2)[:E] + 0.01 * rng.standard_normal((E, 2, 2, 2))
    geom[1] = geom[1][:, ::-1, :]  # one negatively oriented element: det < 0 everywhere, LU solves it, so must we
    Q = int(np.prod([o + 1 for _, o in rules]))
    f = rng.standard_normal((E, Q))
    u, det_fd = fastdiag_project(basis, rules, gb, geom, f)
    _, det, _ = orc.geometry(gb, rules, geom.reshape(E, 2, 4), 2, want_coords=False, want_jinv=False)
    assert np.abs(det_fd - det).max() < 1e-14 * np.abs(det).max()  # det J is linear in the reference coordinates
    M = orc.mass(basis, basis, rules, E, det=det)
    b = orc.l2_dual(basis, rules, f, det=det)
    ref = np.linalg.solve(M, b[..., None])[..., 0]
    err = np.abs(u - ref).max(axis=1) / np.abs(ref).max(axis=1)
    assert err.max() < 1e-11, err
