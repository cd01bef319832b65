"""Drive the UNMODIFIED compiled reference (oracle/_ref) element by element.  TEST INFRASTRUCTURE.

Each function loops the reference's own public API (``interplib.FunctionSpace``,
``DegreesOfFreedom``, ``CoordinateMap``, ``SpaceMap``, ``compute_mass_matrix``,
``projection_l2_dual`` ...) over a batch of elements and stacks the per-element answers into the
batched layouts of ``include/interp_b200.h``.  Used to validate the C oracle, to generate
``tests/golden/`` and as the ``--impl reference`` arm of ``bench.py``.
"""
from __future__ import annotations

import numpy as np

from . import build_ref


def ref():
    return build_ref.import_reference()


def available() -> bool:
    return build_ref.built() or build_ref.available()


def _fs(ip, basis):
    return ip.FunctionSpace(*(ip.BasisSpecs(t, o) for t, o in basis))


def _ispace(ip, rules):
    return ip.IntegrationSpace(*(ip.IntegrationSpecs(o, m) for m, o in rules))


def rule(method, order):
    ip = ref()
    s = ip.IntegrationSpecs(order, method)
    return np.array(s.nodes()), np.array(s.weights())


def basis_table(basis, rule_):
    """B[n][m], D[n][m] via reconstruct of unit DoF vectors would lose bits; use values_at_integration_nodes
    for B and reconstruct_derivative of unit vectors for D (a single product 1.0*D is exact)."""
    ip = ref()
    (bt, bo), (rm, ro) = basis, rule_
    fs = _fs(ip, [(bt, bo)])
    isp = _ispace(ip, [(rm, ro)])
    B = np.array(fs.values_at_integration_nodes(isp)).T.copy()  # [m][n] -> [n][m]
    D = np.empty_like(B)
    for i in range(bo + 1):
        v = np.zeros(bo + 1)
        v[i] = 1.0
        dof = ip.DegreesOfFreedom(fs, v)
        # sum_i D[i][q]*dof[i] with one non-zero unit entry: 0 + D*1 (+0 ...) is exact
        D[i, :] = dof.reconstruct_derivative_at_integration_points(isp, idim=[0])
    return B, D


def space_weights(rules):
    ip = ref()
    return np.array(_ispace(ip, rules).weights())


def reconstruct(basis, rules, dofs, deriv_masks=(0,)):
    ip = ref()
    fs = _fs(ip, basis)
    isp = _ispace(ip, rules)
    shape = [o + 1 for _, o in basis]
    N = int(np.prod(shape))
    Q = int(np.prod([o + 1 for _, o in rules]))
    dofs = np.asarray(dofs, dtype=np.float64).reshape(-1, N)
    out = np.empty((dofs.shape[0], len(deriv_masks), Q))
    d = len(basis)
    for e in range(dofs.shape[0]):
        dof = ip.DegreesOfFreedom(fs, dofs[e].reshape(shape))
        for f, mask in enumerate(deriv_masks):
            if mask == 0:
                v = dof.reconstruct_at_integration_points(isp)
            else:
                v = dof.reconstruct_derivative_at_integration_points(isp, idim=[k for k in range(d) if (mask >> k) & 1])
            out[e, f, :] = np.asarray(v).reshape(-1)
    return out


def space_map(ip, geom_basis, isp, gdofs):
    """gdofs[n_phys][Ng] -> reference SpaceMap"""
    fs = _fs(ip, geom_basis)
    shape = [o + 1 for _, o in geom_basis]
    maps = [ip.CoordinateMap(ip.DegreesOfFreedom(fs, g.reshape(shape)), isp) for g in gdofs]
    return ip._interp.SpaceMap(*maps), maps


def geometry(geom_basis, rules, geom_dofs, n_phys):
    ip = ref()
    isp = _ispace(ip, rules)
    d = len(geom_basis)
    Ng = int(np.prod([o + 1 for _, o in geom_basis]))
    Q = int(np.prod([o + 1 for _, o in rules]))
    g = np.asarray(geom_dofs, dtype=np.float64).reshape(-1, n_phys, Ng)
    E = g.shape[0]
    coords = np.empty((E, n_phys, 1 + d, Q))
    det = np.empty((E, Q))
    jinv = np.empty((E, Q, d, n_phys))
    for e in range(E):
        smap, maps = space_map(ip, geom_basis, isp, g[e])
        for a, m in enumerate(maps):
            coords[e, a, 0] = np.asarray(m.values).reshape(-1)
            for k in range(d):
                coords[e, a, 1 + k] = np.asarray(m.gradient(k)).reshape(-1)
        det[e] = np.asarray(smap.determinant).reshape(-1)
        jinv[e] = np.asarray(smap.inverse_map).reshape(Q, d, n_phys)
    return coords, det, jinv


def mass(basis_in, basis_out, rules, E, geom_basis=None, geom_dofs=None, n_phys=0, idx_in=-1, idx_out=0):
    """Reference mass / gradient-mass matrices, out[E][N_out][N_in].
    geom_dofs None -> IntegrationSpace branch (E identical matrices)."""
    ip = ref()
    fsi, fso = _fs(ip, basis_in), _fs(ip, basis_out)
    isp = _ispace(ip, rules)
    Nin = int(np.prod([o + 1 for _, o in basis_in]))
    Nout = int(np.prod([o + 1 for _, o in basis_out]))
    out = np.empty((E, Nout, Nin))
    if geom_dofs is not None:
        Ng = int(np.prod([o + 1 for _, o in geom_basis]))
        g = np.asarray(geom_dofs, dtype=np.float64).reshape(E, n_phys, Ng)
    for e in range(E):
        integ = isp if geom_dofs is None else space_map(ip, geom_basis, isp, g[e])[0]
        if idx_in < 0:
            out[e] = ip.compute_mass_matrix(fsi, fso, integ)
        else:
            out[e] = ip._interp.compute_gradient_mass_matrix(fsi, fso, integ, idx_in, idx_out)
    return out


def l2_dual(basis, rules, fvals, geom_basis=None, geom_dofs=None, n_phys=0):
    """projection_l2_dual of tabulated samples: the callable handed to the reference returns the
    element's row of fvals (shaped like the node arrays)."""
    ip = ref()
    fs = _fs(ip, basis)
    isp = _ispace(ip, rules)
    N = int(np.prod([o + 1 for _, o in basis]))
    mshape = [o + 1 for _, o in rules]
    Q = int(np.prod(mshape))
    f = np.asarray(fvals, dtype=np.float64).reshape(-1, Q)
    E = f.shape[0]
    out = np.empty((E, N))
    if geom_dofs is not None:
        Ng = int(np.prod([o + 1 for _, o in geom_basis]))
        g = np.asarray(geom_dofs, dtype=np.float64).reshape(E, n_phys, Ng)
    for e in range(E):
        integ = isp if geom_dofs is None else space_map(ip, geom_basis, isp, g[e])[0]
        fe = f[e].reshape(mshape)
        dual = ip.projection_l2_dual(lambda *x, fe=fe: fe, fs, integ)
        out[e] = np.asarray(dual.values).reshape(-1)
    return out


def l2_primal(basis, rules, fvals, geom_basis=None, geom_dofs=None, n_phys=0):
    """The reference's projection_l2_primal (python/interplib/integration.py:171-223: dual DoFs, mass matrix,
    np.linalg.solve) of tabulated samples, element by element -> primal DoFs out[E][N]."""
    ip = ref()
    fs = _fs(ip, basis)
    isp = _ispace(ip, rules)
    N = int(np.prod([o + 1 for _, o in basis]))
    mshape = [o + 1 for _, o in rules]
    Q = int(np.prod(mshape))
    f = np.asarray(fvals, dtype=np.float64).reshape(-1, Q)
    E = f.shape[0]
    out = np.empty((E, N))
    if geom_dofs is not None:
        Ng = int(np.prod([o + 1 for _, o in geom_basis]))
        g = np.asarray(geom_dofs, dtype=np.float64).reshape(E, n_phys, Ng)
    for e in range(E):
        integ = isp if geom_dofs is None else space_map(ip, geom_basis, isp, g[e])[0]
        fe = f[e].reshape(mshape)
        primal = ip.projection_l2_primal(lambda *x, fe=fe: fe, fs, integ)
        out[e] = np.asarray(primal.values).reshape(-1)
    return out


def evaluate_points(basis, x):
    """FunctionSpace.evaluate per element: x[E][ndim][P] -> [E][P][N]"""
    ip = ref()
    fs = _fs(ip, basis)
    x = np.ascontiguousarray(x, dtype=np.float64)
    E, ndim, P = x.shape
    N = int(np.prod([o + 1 for _, o in basis]))
    out = np.empty((E, P, N))
    for e in range(E):
        out[e] = np.asarray(fs.evaluate(*(np.ascontiguousarray(x[e, k]) for k in range(ndim)))).reshape(P, N)
    return out


def reconstruct_points(basis, dofs, x):
    """interplib.reconstruct(dof, *x) per element: dofs[E][N], x[E][ndim][P] -> [E][P]"""
    ip = ref()
    fs = _fs(ip, basis)
    shape = [o + 1 for _, o in basis]
    x = np.ascontiguousarray(x, dtype=np.float64)
    E, ndim, P = x.shape
    dofs = np.asarray(dofs, dtype=np.float64).reshape(E, -1)
    out = np.empty((E, P))
    for e in range(E):
        dof = ip.DegreesOfFreedom(fs, dofs[e].reshape(shape))
        out[e] = ip.reconstruct(dof, *(np.ascontiguousarray(x[e, k]) for k in range(ndim)))
    return out


def basis_points_1d(basis1, x):
    """BasisSpecs.values / .derivatives at the points x[P] -> ([P][n], [P][n])"""
    ip = ref()
    b = ip.BasisSpecs(*basis1)
    x = np.ascontiguousarray(x, dtype=np.float64)
    return np.asarray(b.values(x)).reshape(len(x), -1), np.asarray(b.derivatives(x)).reshape(len(x), -1)


def incidence_matrix(basis1):
    ip = ref()
    from interplib import _interp  # not re-exported by the package (tests/test_incidence.py:9-13 import it from here)

    return np.asarray(_interp.incidence_matrix(ip.BasisSpecs(*basis1)))


def dof_derivative(basis, dofs, idim):
    """DegreesOfFreedom.derivative(idim) per element: dofs[E][N] -> [E][N_lower]"""
    ip = ref()
    fs = _fs(ip, basis)
    shape = [o + 1 for _, o in basis]
    dofs = np.asarray(dofs, dtype=np.float64).reshape(-1, int(np.prod(shape)))
    out = [np.asarray(ip.DegreesOfFreedom(fs, dofs[e].reshape(shape)).derivative(idim).values).reshape(-1)
           for e in range(dofs.shape[0])]
    return np.stack(out)


def covector_table(n):
    """Exhaustive integer table of the reference's CovectorBasis in n dimensions (N4): for every pair of component sets
    (c1, c2): the wedge's (is_zero, sign, rank), for every c1 its Hodge's sign, and the order relation of equal-rank pairs."""
    from itertools import combinations

    ref()  # puts oracle/_ref on sys.path
    from interplib import _interp

    sets = [c for k in range(n + 1) for c in combinations(range(n), k)]
    full = _interp.CovectorBasis(n, *range(n))
    wedge, hodge, order = [], [], []
    for c1 in sets:
        b1 = _interp.CovectorBasis(n, *c1)
        h = ~b1
        # the complement's components are determined; its sign is what has to match: b1 ^ ~b1 == +-full
        hodge.append([h.rank, h.sign, (b1 ^ h).sign if (b1 ^ h) else 0, 1 if (b1 ^ h) and abs((b1 ^ h).sign) == 1 and (b1 ^ h).rank == full.rank else 0])
        for c2 in sets:
            b2 = _interp.CovectorBasis(n, *c2)
            w = b1 ^ b2
            wedge.append([1 if w else 0, w.sign if w else 0, w.rank if w else 0])
            if len(c1) == len(c2):
                order.append([int(b1 < b2), int(b1 == b2), int(b1 > b2)])
    return dict(sets=[list(c) for c in sets], wedge=np.array(wedge), hodge=np.array(hodge), order=np.array(order))
