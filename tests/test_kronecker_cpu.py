"""CPU: the identities the Kronecker-sum mass kernels rely on (csrc/kernels_mass2d_kron.cu, kernels_mass3d_kron.cu, the 1-D
segment kernel in kernels_fastdiag.cu), stated in NumPy with the library's own host tables (ib200_pair_moments) and checked
against the oracle's quadrature loops (restatement of src/python/mass_matrices.c:161-565 and mappings.c:304-556):

* bilinear quad:        det J = alpha + beta xi0 + gamma xi1          ->  M = alpha A0xA1 + beta X0xA1 + gamma A0xX1
* its adjugate:         det J * Jinv[a][b] = a + b xi_(1-a)           ->  gradient matrices, the reference's mirroring included
* trilinear hexahedron: det J = sum_abc d_abc xi0^a xi1^b xi2^c        ->  27-term Kronecker sum
* straight segment:     det J constant                                 ->  M = det A
"""
import itertools

import numpy as np
import pytest

from interplib_b200 import batch

GLL, LU = "lagrange-gauss-lobatto", "lagrange-uniform"


def linear_basis(gb):
    v = batch.basis_at_points(gb, np.array([0.0, 1.0]))
    return v[0], v[1] - v[0]  # phi_i(xi) = a_i + s_i xi


def quad_coefficients(gb, geom):
    """x_xi0, x_xi1, x_xi0xi1 (and the same for y) of bilinear maps from their geometry DoFs [E][2][4]."""
    (a0, s0), (a1, s1) = (linear_basis(b) for b in gb)
    G = np.stack([np.outer(s0, a1).ravel(), np.outer(a0, s1).ravel(), np.outer(s0, s1).ravel()])
    d = np.einsum("kc,epc->epk", G, geom)
    return d[:, 0], d[:, 1]


def jittered_quads(E, rng):
    base = np.array([[0.0, 0.0, 1.0, 1.0], [0.0, 1.0, 0.0, 1.0]])  # [coordinate][i0 i1]
    return (base[None] + rng.uniform(-0.2, 0.2, (E, 2, 4))) * rng.uniform(0.5, 2.0, (E, 1, 1)) + rng.uniform(-3, 3, (E, 2, 1))


@pytest.mark.parametrize("bin_,bout,rules", [
    ([(GLL, 4)] * 2, [(GLL, 4)] * 2, [("gauss", 5)] * 2),
    ([(GLL, 7), (GLL, 8)], [(GLL, 8), (GLL, 7)], [("gauss", 9)] * 2),
    ([("legendre", 5), ("bernstein", 3)], [("lagrange-chebyshev-gauss", 2), ("lagrange-gauss", 6)], [("gauss", 7), ("gauss-lobatto", 8)]),
])
@pytest.mark.parametrize("gb", [[(LU, 1)] * 2, [("legendre", 1), ("lagrange-gauss", 1)]])
def test_bilinear_quad_mass_is_a_kronecker_sum(orc, bin_, bout, rules, gb):
    rng = np.random.default_rng(3)
    E = 6
    corners = jittered_quads(E, rng).reshape(E, 2, 2, 2)
    inv = [np.linalg.inv(batch.basis_at_points(b, np.array([-1.0, 1.0]))) for b in gb]
    geom = np.ascontiguousarray(np.einsum("ia,jb,epab->epij", inv[0], inv[1], corners).reshape(E, 2, 4))
    dx, dy = quad_coefficients(gb, geom)
    alpha = dx[:, 0] * dy[:, 1] - dx[:, 1] * dy[:, 0]
    beta = dx[:, 0] * dy[:, 2] - dx[:, 2] * dy[:, 0]
    gamma = dx[:, 2] * dy[:, 1] - dx[:, 1] * dy[:, 2]
    (A0, X0), (A1, X1) = (batch.pair_moments(bout[k], bin_[k], rules[k], 1) for k in range(2))
    M = (np.einsum("e,ab,cd->eacbd", alpha, A0, A1) + np.einsum("e,ab,cd->eacbd", beta, X0, A1)
         + np.einsum("e,ab,cd->eacbd", gamma, A0, X1)).reshape(E, A0.shape[0] * A1.shape[0], -1)
    _, det, jinv = orc.geometry(gb, rules, geom, 2)
    ref = orc.mass(bin_, bout, rules, E, det=det)
    assert (np.abs(M - ref).reshape(E, -1).max(axis=1) <= 1e-13 * np.abs(ref).reshape(E, -1).max(axis=1)).all()

    # gradient matrices: the adjugate entries are linear in one coordinate; matching spaces are mirrored from the lower triangle
    refs = {(ii, io): orc.mass(bin_, bout, rules, E, det=det, jinv=jinv, n_phys=2, deriv_in=ii, idx_out=io) for ii in range(2) for io in range(2)}
    scale = np.max([np.abs(r).reshape(E, -1).max(axis=1) for r in refs.values()], axis=0)
    for (ii, io), ref in refs.items():
        a, b = {(0, 0): (dy[:, 1], dy[:, 2]), (0, 1): (-dx[:, 1], -dx[:, 2]), (1, 0): (-dy[:, 0], -dy[:, 2]), (1, 1): (dx[:, 0], dx[:, 2])}[(ii, io)]
        T0 = batch.pair_moments(bout[0], bin_[0], rules[0], 1, deriv_in=ii == 0)
        T1 = batch.pair_moments(bout[1], bin_[1], rules[1], 1, deriv_in=ii == 1)
        if ii == 0:
            Mg = np.einsum("e,ab,cd->eacbd", a, T0[0], T1[0]) + np.einsum("e,ab,cd->eacbd", b, T0[1], T1[0])
        else:
            Mg = np.einsum("e,ab,cd->eacbd", a, T0[0], T1[0]) + np.einsum("e,ab,cd->eacbd", b, T0[0], T1[1])
        Mg = Mg.reshape(ref.shape)
        if bin_ == bout:  # mass_matrices.c:497-560: entries on and below the diagonal, copied into the upper triangle
            Mg = np.tril(Mg) + np.swapaxes(np.tril(Mg, -1), 1, 2)
        assert (np.abs(Mg - ref).reshape(E, -1).max(axis=1) <= 1e-12 * scale).all(), (ii, io)


@pytest.mark.parametrize("bin_,bout,rules,gb", [
    ([(GLL, 3)] * 3, [(GLL, 3)] * 3, [("gauss", 4)] * 3, [(LU, 1)] * 3),
    ([(GLL, 2)] * 3, [("legendre", 3), ("bernstein", 1), ("lagrange-gauss", 2)], [("gauss", 4)] * 3, [("bernstein", 1), ("legendre", 1), (LU, 1)]),
])
def test_trilinear_hexahedron_mass_is_a_27_term_kronecker_sum(orc, bin_, bout, rules, gb):
    rng = np.random.default_rng(4)
    E = 4
    corners = np.array([[[[2 * i0 - 1, 2 * i1 - 1, 2 * i2 - 1] for i2 in range(2)] for i1 in range(2)] for i0 in range(2)], dtype=float)
    corners = (corners[None] + rng.uniform(-0.2, 0.2, (E, 2, 2, 2, 3))).transpose(0, 4, 1, 2, 3)
    inv = [np.linalg.inv(batch.basis_at_points(b, np.array([-1.0, 1.0]))) for b in gb]
    geom = np.ascontiguousarray(np.einsum("ia,jb,kc,epabc->epijk", inv[0], inv[1], inv[2], corners).reshape(E, 3, 8))
    m = [np.stack(linear_basis(b)) for b in gb]  # [dim][power][i]
    T = np.einsum("ai,bj,ck->abcijk", m[0], m[1], m[2]).reshape(8, 8)
    Xm = np.einsum("mn,epn->epm", T, geom).reshape(E, 3, 2, 2, 2)
    d = np.zeros((E, 3, 3, 3))
    perms = [((0, 1, 2), 1), ((1, 2, 0), 1), ((2, 0, 1), 1), ((0, 2, 1), -1), ((2, 1, 0), -1), ((1, 0, 2), -1)]
    for (p0, p1, p2), sg in perms:
        for b0, c0, a1, c1, a2, b2 in itertools.product(range(2), repeat=6):
            d[:, a1 + a2, b0 + b2, c0 + c1] += sg * Xm[:, p0, 1, b0, c0] * Xm[:, p1, a1, 1, c1] * Xm[:, p2, a2, b2, 1]
    mom = [batch.pair_moments(bout[k], bin_[k], rules[k], 2) for k in range(3)]
    M = np.einsum("eabc,ajk,blm,cno->ejlnkmo", d, mom[0], mom[1], mom[2])
    M = M.reshape(E, int(np.prod([x.shape[1] for x in mom])), -1)
    _, det, _ = orc.geometry(gb, rules, geom, 3, want_coords=False, want_jinv=False)
    ref = orc.mass(bin_, bout, rules, E, det=det)
    assert (np.abs(M - ref).reshape(E, -1).max(axis=1) <= 1e-13 * np.abs(ref).reshape(E, -1).max(axis=1)).all()


def test_straight_segment_mass_is_det_times_the_reference_matrix(orc):
    rng = np.random.default_rng(5)
    E, basis, rule, gb = 12, ("legendre", 5), ("gauss-lobatto", 7), ("bernstein", 1)
    knots = np.sort(rng.random((E, 2)), axis=1) + np.array([0.0, 0.3])
    knots[2] = knots[2][::-1]
    g = np.ascontiguousarray((knots @ np.linalg.inv(batch.basis_at_points(gb, np.array([-1.0, 1.0]))).T).reshape(E, 1, 2))
    a, s = linear_basis(gb)
    det1 = g[:, 0, :] @ s
    _, det, _ = orc.geometry([gb], [rule], g, 1, want_coords=False, want_jinv=False)
    assert np.allclose(det, det1[:, None], rtol=1e-14)
    A = batch.pair_moments(basis, basis, rule, 0)[0]
    ref = orc.mass([basis], [basis], [rule], E, det=det)
    assert np.abs(det1[:, None, None] * A - ref).max() <= 1e-14 * np.abs(ref).max()
