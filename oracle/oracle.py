"""ctypes front-end of the CPU oracle (``oracle/interp_oracle.c``).  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import this module; the product (``interplib_b200``) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
SRC = HERE / "interp_oracle.c"
LIB = HERE / "_build" / "liboracle.so"

BASIS_TYPES = {
    "legendre": 1,
    "lagrange-gauss-lobatto": 2,
    "lagrange-gauss": 3,
    "lagrange-uniform": 4,
    "lagrange-chebyshev-gauss": 5,
    "bernstein": 6,
}
RULE_TYPES = {"gauss": 1, "gauss-lobatto": 2}


def build(force: bool = False) -> Path:
    """Compile the oracle with gcc (no -march: no FMA contraction, plain IEEE doubles)."""
    if LIB.exists() and not force and (not SRC.exists() or LIB.stat().st_mtime > SRC.stat().st_mtime):
        return LIB
    LIB.parent.mkdir(parents=True, exist_ok=True)
    cmd = ["gcc", "-std=gnu17", "-O2", "-fPIC", "-ffp-contract=off", "-fopenmp", "-shared", str(SRC), "-o", str(LIB), "-lm"]
    subprocess.run(cmd, check=True)
    return LIB


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(build()))
    return _lib


def _ints(vals):
    return (C.c_int * len(vals))(*[int(v) for v in vals])


def _uints(vals):
    return (C.c_uint * len(vals))(*[int(v) for v in vals])


def _dp(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _split(specs, table):
    """[(type_name, order), ...] -> (c_int[], c_uint[])"""
    return _ints([table[t] for t, _ in specs]), _uints([o for _, o in specs])


def _chk(res, what):
    if res != 0:
        raise RuntimeError(f"oracle {what} failed with code {res}")


def threads() -> int:
    return int(lib().orc_threads())


def set_threads(n: int) -> None:
    os.environ["OMP_NUM_THREADS"] = str(n)


def rule(method: str, order: int):
    x = np.empty(order + 1)
    w = np.empty(order + 1)
    _chk(lib().orc_rule(RULE_TYPES[method], C.c_uint(order), _dp(x), _dp(w)), "rule")
    return x, w


def roots(btype: str, order: int):
    r = np.empty(order + 1)
    _chk(lib().orc_roots(BASIS_TYPES[btype], C.c_uint(order), _dp(r)), "roots")
    return r


def basis_table(basis, rule_):
    (bt, bo), (rt, ro) = basis, rule_
    B = np.empty((bo + 1, ro + 1))
    D = np.empty((bo + 1, ro + 1))
    _chk(lib().orc_basis_table(BASIS_TYPES[bt], C.c_uint(bo), RULE_TYPES[rt], C.c_uint(ro), _dp(B), _dp(D)), "basis_table")
    return B, D


def _counts(basis, rules):
    N = int(np.prod([o + 1 for _, o in basis]))
    Q = int(np.prod([o + 1 for _, o in rules]))
    return N, Q


def space_weights(rules):
    rt, ro = _split(rules, RULE_TYPES)
    W = np.empty([o + 1 for _, o in rules])
    _chk(lib().orc_space_weights(C.c_uint(len(rules)), rt, ro, _dp(W)), "space_weights")
    return W


def reconstruct(basis, rules, dofs, deriv_masks=(0,)):
    """dofs[E][N] -> out[E][n_fields][Q]"""
    N, Q = _counts(basis, rules)
    dofs = np.ascontiguousarray(dofs, dtype=np.float64).reshape(-1, N)
    E = dofs.shape[0]
    out = np.empty((E, len(deriv_masks), Q))
    bt, bo = _split(basis, BASIS_TYPES)
    rt, ro = _split(rules, RULE_TYPES)
    _chk(
        lib().orc_reconstruct(C.c_uint(len(basis)), bt, bo, rt, ro, C.c_size_t(E), _dp(dofs), C.c_uint(len(deriv_masks)),
                              _uints(deriv_masks), _dp(out)),
        "reconstruct",
    )
    return out


def evaluate_points(basis, x):
    """x[E][ndim][P] -> out[E][P][N]   (FunctionSpace.evaluate per element)"""
    x = np.ascontiguousarray(x, dtype=np.float64)
    E, ndim, P = x.shape
    N = int(np.prod([o + 1 for _, o in basis]))
    out = np.empty((E, P, N))
    bt, bo = _split(basis, BASIS_TYPES)
    _chk(lib().orc_evaluate_points(C.c_uint(ndim), bt, bo, C.c_size_t(E), C.c_size_t(P), _dp(x), _dp(out)), "evaluate_points")
    return out


def reconstruct_points(basis, dofs, x, deriv_masks=(0,)):
    """dofs[E][N], x[E][ndim][P] -> out[E][n_fields][P]   (reconstruct() per element)"""
    x = np.ascontiguousarray(x, dtype=np.float64)
    E, ndim, P = x.shape
    N = int(np.prod([o + 1 for _, o in basis]))
    dofs = np.ascontiguousarray(dofs, dtype=np.float64).reshape(E, N)
    out = np.empty((E, len(deriv_masks), P))
    bt, bo = _split(basis, BASIS_TYPES)
    _chk(
        lib().orc_reconstruct_points(C.c_uint(ndim), bt, bo, C.c_size_t(E), _dp(dofs), C.c_size_t(P), _dp(x),
                                     C.c_uint(len(deriv_masks)), _uints(deriv_masks), _dp(out)),
        "reconstruct_points",
    )
    return out


def incidence_matrix(basis1):
    """[n][n+1] (incidence_matrix of a 1-D basis)"""
    (bt,), (bo,) = [BASIS_TYPES[basis1[0]]], [basis1[1]]
    out = np.empty((bo, bo + 1))
    _chk(lib().orc_incidence_matrix(C.c_int(bt), C.c_uint(bo), _dp(out)), "incidence_matrix")
    return out


def dof_derivative(basis, dofs, idim):
    """dofs[E][N] -> [E][N_lower]   (DegreesOfFreedom.derivative per element)"""
    N = int(np.prod([o + 1 for _, o in basis]))
    Nl = int(np.prod([o + (0 if k == idim else 1) for k, (_, o) in enumerate(basis)]))
    dofs = np.ascontiguousarray(dofs, dtype=np.float64).reshape(-1, N)
    E = dofs.shape[0]
    out = np.empty((E, Nl))
    bt, bo = _split(basis, BASIS_TYPES)
    _chk(lib().orc_dof_derivative(C.c_uint(len(basis)), bt, bo, C.c_uint(idim), C.c_size_t(E), _dp(dofs), _dp(out)),
         "dof_derivative")
    return out


def geometry(geom_basis, rules, geom_dofs, n_phys, want_coords=True, want_det=True, want_jinv=True):
    """geom_dofs[E][n_phys][Ng] -> coords[E][n_phys][1+d][Q], det[E][Q], jinv[E][Q][d][n_phys]"""
    Ng, Q = _counts(geom_basis, rules)
    d = len(geom_basis)
    g = np.ascontiguousarray(geom_dofs, dtype=np.float64).reshape(-1, n_phys, Ng)
    E = g.shape[0]
    coords = np.empty((E, n_phys, 1 + d, Q)) if want_coords else None
    det = np.empty((E, Q)) if want_det else None
    jinv = np.empty((E, Q, d, n_phys)) if want_jinv else None
    bt, bo = _split(geom_basis, BASIS_TYPES)
    rt, ro = _split(rules, RULE_TYPES)
    _chk(
        lib().orc_geometry(C.c_uint(d), C.c_uint(n_phys), bt, bo, rt, ro, C.c_size_t(E), _dp(g), _dp(coords), _dp(det), _dp(jinv)),
        "geometry",
    )
    return coords, det, jinv


def mass(basis_in, basis_out, rules, E, det=None, jinv=None, n_phys=0, deriv_in=-1, idx_out=0):
    """-> out[E][N_out][N_in]; det[E][Q] / jinv[E][Q][d][n_phys] optional"""
    Nin, Q = _counts(basis_in, rules)
    Nout, _ = _counts(basis_out, rules)
    out = np.empty((E, Nout, Nin))
    bti, boi = _split(basis_in, BASIS_TYPES)
    bto, boo = _split(basis_out, BASIS_TYPES)
    rt, ro = _split(rules, RULE_TYPES)
    if det is not None:
        det = np.ascontiguousarray(det, dtype=np.float64).reshape(E, Q)
    if jinv is not None:
        jinv = np.ascontiguousarray(jinv, dtype=np.float64)
    _chk(
        lib().orc_mass(C.c_uint(len(rules)), bti, boi, bto, boo, rt, ro, C.c_size_t(E), _dp(det), _dp(jinv), C.c_uint(n_phys),
                       C.c_int(deriv_in), C.c_int(idx_out), _dp(out)),
        "mass",
    )
    return out


def l2_dual(basis, rules, fvals, det=None):
    """fvals[E][Q] (+det[E][Q]) -> out[E][N]"""
    N, Q = _counts(basis, rules)
    f = np.ascontiguousarray(fvals, dtype=np.float64).reshape(-1, Q)
    E = f.shape[0]
    if det is not None:
        det = np.ascontiguousarray(det, dtype=np.float64).reshape(E, Q)
    out = np.empty((E, N))
    bt, bo = _split(basis, BASIS_TYPES)
    rt, ro = _split(rules, RULE_TYPES)
    _chk(lib().orc_l2_dual(C.c_uint(len(rules)), bt, bo, rt, ro, C.c_size_t(E), _dp(f), _dp(det), _dp(out)), "l2_dual")
    return out
