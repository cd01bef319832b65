"""k-form bookkeeping: drop-in mirror of the reference's ``CovectorBasis`` type
(/root/reference/src/python/covector_basis.c) over the C ABI (``ib200_covector_*``, ``ib200_kform_*``), and the glue that
turns a k-form on a tensor-product element into the list of component spaces whose block mass matrices the GPU computes
(``Context.mass_matrix_mapped`` per pair of components -- BASELINE config 3 is the 2-D 1-form of order 8)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import Covector, basis_specs, check, lib


class CovectorBasis:
    """``CovectorBasis(n, *idx, sign=...)``: the wedge product of the ``dx_i`` listed (ascending, unique) in ``n``
    dimensions.  ``^`` wedge, ``~`` Hodge, unary ``-`` sign flip, ordering by the reference's rule
    (covector_basis.c:100-125; comparing bundles of different rank raises ``ValueError``)."""

    __slots__ = ("_c",)

    def __init__(self, n: int, /, *idx: int, sign: int = 1):
        c = Covector()
        arr = (C.c_uint * max(len(idx), 1))(*[int(i) for i in idx])
        if any(int(i) < 0 for i in idx):
            raise ValueError("Component indices must be non-negative.")
        check(lib.ib200_covector_make(int(n), -1 if sign < 0 else 1, len(idx), arr, C.byref(c)))
        self._c = c

    @classmethod
    def _wrap(cls, c: Covector) -> "CovectorBasis":
        obj = object.__new__(cls)
        obj._c = c
        return obj

    ndim = property(lambda self: self._c.dimension)
    rank = property(lambda self: int(lib.ib200_covector_rank(self._c)))
    sign = property(lambda self: -1 if self._c.sign else 1)
    bits = property(lambda self: int(self._c.basis_bits))
    components = property(lambda self: tuple(i for i in range(self.ndim) if self.bits >> i & 1))

    def __bool__(self) -> bool:
        return not lib.ib200_covector_is_zero(self._c)

    def __xor__(self, other: "CovectorBasis") -> "CovectorBasis":
        if not isinstance(other, CovectorBasis):
            return NotImplemented
        if not self or not other:  # covector_basis_xor, covector_basis.c:80-118: a zero operand gives zero
            return CovectorBasis._wrap(Covector())
        if self.ndim != other.ndim:
            raise ValueError(f"Cannot xor two basis sets with different dimensions ({self.ndim} != {other.ndim}).")
        out = Covector()
        check(lib.ib200_covector_wedge(self._c, other._c, C.byref(out)))
        return CovectorBasis._wrap(out)

    def __invert__(self) -> "CovectorBasis":
        out = Covector()
        check(lib.ib200_covector_hodge(self._c, C.byref(out)))
        return CovectorBasis._wrap(out)

    def __neg__(self) -> "CovectorBasis":
        return CovectorBasis._wrap(Covector.make(self._c.dimension, not self._c.sign, self._c.basis_bits))

    def apply_contra(self, dim: int) -> "CovectorBasis":
        out = Covector()
        check(lib.ib200_covector_apply_contra(self._c, int(dim), C.byref(out)))
        return CovectorBasis._wrap(out)

    def _compare(self, other, op: str):
        """covector_basis_richcompare, src/python/covector_basis.c:150-205."""
        if not isinstance(other, CovectorBasis):
            return NotImplemented
        if self._c.sign != other._c.sign:
            if op == "ne":
                return True
            if op == "eq":
                return False
            raise ValueError("Can not compare two basis with different signs for relation other than equal/not equal.")
        if not self and not other:
            return op == "eq"
        rel = int(lib.ib200_covector_order(self._c, other._c))
        if rel == 3:
            raise ValueError(f"Cannot compare two bases of different dimensions and/or ranks (({self.ndim}, {self.rank}) != "
                             f"({other.ndim}, {other.rank})).")
        return op in {1: ("eq", "ge", "le"), 2: ("gt", "ge", "ne"), 0: ("lt", "le", "ne")}[rel]

    def __eq__(self, other):
        return self._compare(other, "eq")

    def __ne__(self, other):
        return self._compare(other, "ne")

    def __lt__(self, other):
        return self._compare(other, "lt")

    def __le__(self, other):
        return self._compare(other, "le")

    def __gt__(self, other):
        return self._compare(other, "gt")

    def __ge__(self, other):
        return self._compare(other, "ge")

    def __hash__(self):
        return hash((self._c.dim_sign, self._c.basis_bits))

    def __repr__(self):
        if not self:
            return "CovectorBasis(0)"
        body = " ^ ".join(f"dx_{i}" for i in self.components) or "1"
        return f"{'-' if self.sign < 0 else '+'}{body} in {self.ndim}D"


def kform_components(ndim: int, k: int) -> list[CovectorBasis]:
    """Components of a k-form in ascending order (ib200_kform_components)."""
    n = int(lib.ib200_kform_component_count(ndim, k))
    arr = (Covector * max(n, 1))()
    check(lib.ib200_kform_components(ndim, k, arr))
    return [CovectorBasis._wrap(Covector(arr[i].dim_sign, arr[i].basis_bits)) for i in range(n)]


def kform_component_space(basis, component: CovectorBasis) -> list[tuple[str, int]]:
    """Function space of one component of a k-form over the space ``basis`` = [(type, order), ...]: the order drops by
    one along every dimension the component contains (ib200_kform_component_space)."""
    d = len(basis)
    out = (_lib.BasisSpec * d)()
    check(lib.ib200_kform_component_space(d, basis_specs(basis), component._c, out))
    names = {v: k for k, v in _lib.BASIS_TYPES.items()}
    return [(names[out[i].type], int(out[i].order)) for i in range(d)]


def kform_spaces(basis, k: int) -> list[list[tuple[str, int]]]:
    """Component spaces of the k-form in component order; the 2-D 1-form of order 8 gives [(7, 8)], [(8, 7)] --
    BASELINE config 3's two spaces."""
    return [kform_component_space(basis, c) for c in kform_components(len(basis), k)]


def kform_dof_count(basis, k: int) -> int:
    n = C.c_size_t()
    check(lib.ib200_kform_dof_count(len(basis), basis_specs(basis), k, C.byref(n)))
    return int(n.value)


def dof_offsets(counts) -> np.ndarray:
    """offsets[E + 1] of discontinuous per-element DoF blocks (ib200_dof_offsets)."""
    counts = np.ascontiguousarray(counts, dtype=np.uintp)
    out = np.empty(counts.size + 1, dtype=np.uintp)
    check(lib.ib200_dof_offsets(counts.size, counts.ctypes.data_as(C.POINTER(C.c_size_t)), out.ctypes.data_as(C.POINTER(C.c_size_t))))
    return out


def kform_mass_blocks(ctx, basis, k: int, rules, geom_basis, geom_dofs, n_phys: int, **kw) -> dict:
    """All blocks M[a][b] = compute_mass_matrix(space_b, space_a, smap) of the determinant-weighted k-form mass matrix,
    one GPU launch per block (symmetric blocks and their transposes included, as the reference would compute them);
    returns {(a, b): array[E][N_a][N_b]} in component order."""
    spaces = kform_spaces(basis, k)
    return {(a, b): ctx.mass_matrix_mapped(sb, sa, rules, geom_basis, geom_dofs, n_phys, **kw)
            for a, sa in enumerate(spaces) for b, sb in enumerate(spaces)}


def kform_mass_matrix(ctx, basis, k: int, rules, geom_basis, geom_dofs, n_phys: int, **kw):
    """The k-form element mass matrices M[E][N][N] in one call (one GPU launch for bilinear quads in the plane); block
    (a, b) of M[e] is what kform_mass_blocks returns under (a, b)."""
    return ctx.kform_mass_mapped(basis, k, rules, geom_basis, geom_dofs, n_phys, kform_dof_count(basis, k), **kw)


def assemble_blocks(blocks: dict) -> np.ndarray:
    """Host-side layout of {(a, b): [E][N_a][N_b]} into one [E][N][N] element matrix (N = kform_dof_count)."""
    n = 1 + max(a for a, _ in blocks)
    return np.concatenate([np.concatenate([blocks[(a, b)] for b in range(n)], axis=2) for a in range(n)], axis=1)
