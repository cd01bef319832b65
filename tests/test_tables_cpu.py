"""CPU: the product's host tables (libinterp_b200.so, rows A1-A4/A10/A11/A13) must be bit-equal to the oracle
and to the compiled reference."""
import numpy as np
import pytest

from interplib_b200 import batch as B

BASIS = ["legendre", "lagrange-gauss-lobatto", "lagrange-gauss", "lagrange-uniform", "lagrange-chebyshev-gauss", "bernstein"]
RULES = ["gauss", "gauss-lobatto"]


@pytest.mark.parametrize("method", RULES)
def test_rules_bit_equal_oracle(orc, method):
    for order in range(0, 33):
        x, w, acc = B.rule_1d(method, order)
        xo, wo = orc.rule(method, order)
        assert np.array_equal(x, xo) and np.array_equal(w, wo), (method, order)
        if method == "gauss":
            assert acc == (2 * order - 1 if order > 0 else 0)
        else:
            assert acc == (2 * order - 3 if order > 1 else order)


def test_rule_quirks():
    # reference quirk (gauss_legendre.c:17-24): the 2-point rules are (-1,+1) with unit weights
    for method in RULES:
        x, w, _ = B.rule_1d(method, 1)
        assert list(x) == [-1.0, 1.0] and list(w) == [1.0, 1.0]
        x, w, _ = B.rule_1d(method, 0)
        assert list(x) == [0.0] and list(w) == [2.0]


@pytest.mark.parametrize("method", RULES)
def test_rule_exactness(method):
    # test_c/integration/test_glg.c, test_gll.c: monomials up to `accuracy` are integrated exactly
    for order in range(2, 20):
        x, w, acc = B.rule_1d(method, order)
        for k in range(acc + 1):
            exact = 0.0 if k % 2 else 2.0 / (k + 1)
            assert abs(np.sum(w * x**k) - exact) < 1e-12


@pytest.mark.parametrize("btype", BASIS)
def test_basis_tables_bit_equal_oracle(orc, btype):
    for bo in range(0, 18):
        if btype == "lagrange-uniform" and bo == 0:
            continue  # node 0/0 = NaN in the reference too; values are 1 regardless
        for method in RULES:
            for ro in (0, 1, 2, 3, 4, 7, 9, 17):
                Bv, Dv = B.basis_table_1d((btype, bo), (method, ro))
                Bo, Do = orc.basis_table((btype, bo), (method, ro))
                assert np.array_equal(Bv, Bo), (btype, bo, method, ro)
                assert np.array_equal(Dv, Do), (btype, bo, method, ro)


@pytest.mark.parametrize("btype", [b for b in BASIS if b.startswith("lagrange")])
def test_roots_bit_equal_oracle(orc, btype):
    for bo in range(1, 20):
        assert np.array_equal(B.lagrange_roots(btype, bo), orc.roots(btype, bo))


def test_lagrange_is_nodal():
    for btype in ("lagrange-gauss-lobatto", "lagrange-gauss"):
        method = "gauss-lobatto" if btype.endswith("lobatto") else "gauss"
        Bv, _ = B.basis_table_1d((btype, 5), (method, 5))
        assert np.allclose(Bv, np.eye(6), atol=1e-9)


def test_space_weights_and_nodes(orc):
    rules = [("gauss", 3), ("gauss-lobatto", 4), ("gauss", 1)]
    W = B.integration_space_weights(rules)
    assert np.array_equal(W, orc.space_weights(rules))
    X = B.integration_space_nodes(rules)
    for k, (m, o) in enumerate(rules):
        x, _, _ = B.rule_1d(m, o)
        shape = [1, 1, 1]
        shape[k] = o + 1
        assert np.array_equal(X[k], np.broadcast_to(x.reshape(shape), X[k].shape))


def test_values_at_nodes_vs_ref(refdrive):
    ip = refdrive.ref()
    basis = [("bernstein", 2), ("lagrange-gauss-lobatto", 3)]
    rules = [("gauss", 4), ("gauss-lobatto", 2)]
    fs = ip.FunctionSpace(*(ip.BasisSpecs(t, o) for t, o in basis))
    isp = ip.IntegrationSpace(*(ip.IntegrationSpecs(o, m) for m, o in rules))
    assert np.array_equal(B.values_at_integration_nodes(basis, rules), fs.values_at_integration_nodes(isp))
    assert np.array_equal(B.integration_space_weights(rules), isp.weights())
    assert np.array_equal(B.integration_space_nodes(rules), isp.nodes())


@pytest.mark.parametrize("btype", BASIS)
def test_tables_bit_equal_reference(refdrive, btype):
    for bo in (1, 2, 4, 8):
        for method in RULES:
            for ro in (0, 1, 2, 5, 9):
                Bv, Dv = B.basis_table_1d((btype, bo), (method, ro))
                Br, Dr = refdrive.basis_table((btype, bo), (method, ro))
                assert np.array_equal(Bv, Br) and np.array_equal(Dv, Dr), (btype, bo, method, ro)


@pytest.mark.parametrize("btype", BASIS)
def test_evaluate_bit_equal_reference(refdrive, btype):
    # FunctionSpace.evaluate / BasisSpecs.values / .derivatives (A13): arbitrary points, exact `==`
    ip = refdrive.ref()
    rng = np.random.default_rng(1512)
    for order in (1, 3, 6):
        for npts in (1, 2, 3, 7):
            x = rng.uniform(-1, 1, npts)
            spec = ip.BasisSpecs(btype, order)
            assert np.array_equal(B.basis_at_points((btype, order), x), spec.values(x)), (btype, order, npts)
            assert np.array_equal(B.basis_at_points((btype, order), x, derivative=True), spec.derivatives(x))
    basis = [(btype, 2), ("legendre", 3)]
    xs = [rng.uniform(-1, 1, (4, 3)), rng.uniform(-1, 1, (4, 3))]
    fs = ip.FunctionSpace(*(ip.BasisSpecs(t, o) for t, o in basis))
    assert np.array_equal(B.function_space_evaluate(basis, xs), fs.evaluate(*xs))
