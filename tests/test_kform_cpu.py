"""CPU: k-form bookkeeping (ib200_covector_* / ib200_kform_* through the C ABI, integers only).
Bit-exact against tests/golden/covector.npz -- exhaustive wedge / Hodge / ordering tables of the compiled reference's
CovectorBasis for 1..5 dimensions (tests/golden/make_golden_covector.py) -- plus the reference's own test properties
(/root/reference/tests/test_covector_basis.py) and the DoF counts of src/basis/de_rham_sequence.h:66-110."""
from itertools import combinations
from pathlib import Path

import numpy as np
import pytest

from interplib_b200 import kform as kf
from interplib_b200.kform import CovectorBasis

GOLD = dict(np.load(Path(__file__).resolve().parent / "golden" / "covector.npz"))


def _table(n):
    sets = [c for k in range(n + 1) for c in combinations(range(n), k)]
    full = CovectorBasis(n, *range(n))
    wedge, hodge, order = [], [], []
    for c1 in sets:
        b1 = CovectorBasis(n, *c1)
        h = ~b1
        w = b1 ^ h
        hodge.append([h.rank, h.sign, w.sign if w else 0, 1 if w and abs(w.sign) == 1 and w.rank == full.rank else 0])
        for c2 in sets:
            b2 = CovectorBasis(n, *c2)
            w = b1 ^ b2
            wedge.append([1 if w else 0, w.sign if w else 0, w.rank if w else 0])
            if len(c1) == len(c2):
                order.append([int(b1 < b2), int(b1 == b2), int(b1 > b2)])
    return np.array(wedge), np.array(hodge), np.array(order)


@pytest.mark.parametrize("n", range(1, 6))
def test_covector_tables_match_reference(n):
    wedge, hodge, order = _table(n)
    assert np.array_equal(wedge, GOLD[f"n{n}_wedge"])
    assert np.array_equal(hodge, GOLD[f"n{n}_hodge"])
    assert np.array_equal(order, GOLD[f"n{n}_order"])


def test_covector_tables_vs_reference_fresh(refdrive):
    t = refdrive.covector_table(6)
    wedge, hodge, order = _table(6)
    assert np.array_equal(wedge, t["wedge"]) and np.array_equal(hodge, t["hodge"]) and np.array_equal(order, t["order"])


def test_1d():
    """tests/test_covector_basis.py:9-20"""
    b00 = CovectorBasis(1)
    b01 = CovectorBasis(1, 0)
    assert ~b01 == b00
    assert b01 == ~b00
    assert b00 ^ b01 == b01
    assert b01 ^ b00 == b01
    assert not b01 ^ b01
    assert b00 ^ b00 == b00


@pytest.mark.parametrize("n", range(2, 5))
def test_nd(n):
    """tests/test_covector_basis.py:23-55"""
    b_full = CovectorBasis(n, *range(n))
    b_empty = CovectorBasis(n)
    for k1 in range(1, n):
        for c1 in combinations(range(n), k1):
            b1 = CovectorBasis(n, *c1)
            assert b1 ^ ~b1 == b_full
            assert (~(~b1)).sign == (-1) ** (k1 * (n - k1))
            assert ~b1 ^ b1 == (-b_full if ((k1 * (n - k1)) % 2) else b_full)
            assert b1 ^ b_empty == b1
            assert not (b1 ^ b1)
            for k2 in range(1, n):
                for c2 in combinations(range(n), k2):
                    b2 = CovectorBasis(n, *c2)
                    if k1 != k2:
                        with pytest.raises(ValueError):
                            _ = b1 == b2
                    else:
                        if len(c1) <= n // 2:
                            v1, v2 = c1, c2
                        else:
                            v1 = tuple(i for i in range(n) if i not in c1)
                            v2 = tuple(i for i in range(n) if i not in c2)
                        assert (b1 < b2) == (v1 < v2)
                        assert (b1 <= b2) == (v1 <= v2)
                        assert (b1 >= b2) == (v1 >= v2)
                        assert (b1 > b2) == (v1 > v2)
                        assert (b1 == b2) == (v1 == v2)
                        assert (b1 != b2) == (v1 != v2)


def test_bad_arguments():
    with pytest.raises(ValueError):
        CovectorBasis(3, 2, 1)  # not ascending
    with pytest.raises(ValueError):
        CovectorBasis(2, 2)  # out of range
    with pytest.raises(ValueError):
        CovectorBasis(2, 0) ^ CovectorBasis(3, 0)
    with pytest.raises(ValueError):
        _ = CovectorBasis(2, 0) < -CovectorBasis(2, 1)  # different signs only compare for (in)equality
    assert CovectorBasis(2, 0) != -CovectorBasis(2, 0)


def test_apply_contra_basis():
    """covector_basis_apply_contra_basis (src/basis/covector_basis.c:89-99; test_c/basis/test_covector_basis.c): removing
    dx_d and wedging it back from the left restores the bundle."""
    for n in range(1, 5):
        for k in range(1, n + 1):
            for c in combinations(range(n), k):
                b = CovectorBasis(n, *c)
                for d in c:
                    r = b.apply_contra(d)
                    assert r.components == tuple(i for i in c if i != d) and r.rank == k - 1
                with pytest.raises(ValueError):
                    b.apply_contra(next(i for i in range(n + 1) if i not in c))


def test_kform_components_spaces_and_counts():
    p = 8
    basis = [("lagrange-gauss-lobatto", p)] * 2
    comps = kf.kform_components(2, 1)
    assert [c.components for c in comps] == [(0,), (1,)]
    assert all(a < b for a, b in zip(comps, comps[1:]))
    # BASELINE config 3: the two spaces of the 2-D 1-form of order 8 and its 72 + 72 DoFs
    assert kf.kform_spaces(basis, 1) == [[("lagrange-gauss-lobatto", 7), ("lagrange-gauss-lobatto", 8)],
                                         [("lagrange-gauss-lobatto", 8), ("lagrange-gauss-lobatto", 7)]]
    from interplib_b200 import synthetic as syn

    assert [tuple(map(tuple, s)) for s in kf.kform_spaces(basis, 1)] == [tuple(s) for s in syn.C3_SPACES]
    # differential_form_degrees_of_freedom_count, src/basis/de_rham_sequence.h:66-110
    for o0, o1 in [(1, 1), (3, 5), (8, 8)]:
        b = [("legendre", o0), ("legendre", o1)]
        assert kf.kform_dof_count(b, 0) == (o0 + 1) * (o1 + 1)
        assert kf.kform_dof_count(b, 1) == (o0 + 1) * o1 + o0 * (o1 + 1)
        assert kf.kform_dof_count(b, 2) == o0 * o1
        assert kf.kform_dof_count(b[:1], 0) == o0 + 1 and kf.kform_dof_count(b[:1], 1) == o0
    # components of every rank are sorted and complete
    for n in range(1, 6):
        for k in range(n + 1):
            cs = kf.kform_components(n, k)
            assert len({c.bits for c in cs}) == len(list(combinations(range(n), k)))
            assert all(a < b for a, b in zip(cs, cs[1:]))
    # element_collection_form_offsets, src/basis/de_rham_sequence.c:62-88
    counts = [kf.kform_dof_count([("legendre", p0), ("legendre", p1)], 1) for p0, p1 in [(2, 2), (3, 1), (8, 8)]]
    assert kf.dof_offsets(counts).tolist() == [0, counts[0], counts[0] + counts[1], sum(counts)]
    assert kf.dof_offsets([]).tolist() == [0]
