"""Host-side half of the drop-in mirror (interplib_b200.interp): the properties the reference's own
tests/test_integration_rule.py, test_integration_space.py and test_function_space.py pin, restated against the
mirror.  Everything here runs on the C host-table layer of libinterp_b200.so -- no GPU, no oracle."""
import itertools

import numpy as np
import pytest

from interplib_b200.interp import (BasisRegistry, BasisSpecs, BasisType, DegreesOfFreedom, FunctionSpace, IntegrationMethod,
                                   IntegrationRegistry, IntegrationSpace, IntegrationSpecs, Line, Quad, compute_mass_matrix,
                                   reconstruct)

ALL_TYPES = list(BasisType)


@pytest.mark.parametrize("method", list(IntegrationMethod))
@pytest.mark.parametrize("order", range(1, 12))
def test_rule_integrates_monomials_exactly(method, order):
    # reference tests/test_integration_rule.py:16-28 and test_c/integration/test_glg.c, test_gll.c
    specs = IntegrationSpecs(order, method)
    if order == 1 and method == IntegrationMethod.GAUSS:
        # reference quirk kept for parity: the 2-point "Gauss" rule has nodes -1, +1 (gauss_legendre.c:17-24)
        assert np.array_equal(specs.nodes(), [-1.0, 1.0]) and np.array_equal(specs.weights(), [1.0, 1.0])
        return
    x, w = specs.nodes(), specs.weights()
    assert x.shape == w.shape == (order + 1,) and np.all(np.diff(x) > 0)
    for k in range(specs.accuracy + 1):
        exact = 0.0 if k % 2 else 2.0 / (k + 1)
        assert np.sum(w * x**k) == pytest.approx(exact, abs=1e-13)


def test_specs_validation_and_identity():
    with pytest.raises(ValueError):
        IntegrationSpecs(3, "simpson")
    with pytest.raises(ValueError):
        IntegrationSpecs(-1)
    with pytest.raises(ValueError):
        BasisSpecs("fourier", 2)
    with pytest.raises(TypeError):
        IntegrationSpace()
    with pytest.raises(TypeError):
        FunctionSpace(IntegrationSpecs(2))
    assert IntegrationSpecs(3).method == "gauss" and IntegrationSpecs(3, "gauss-lobatto").order == 3
    assert BasisSpecs(BasisType.LEGENDRE, 4) == BasisSpecs("legendre", 4) != BasisSpecs("legendre", 3)
    fs = FunctionSpace(BasisSpecs("bernstein", 2), BasisSpecs("legendre", 3))
    assert fs.dimension == 2 and fs.orders == (2, 3) and fs == FunctionSpace(*fs.basis_specs)
    assert fs.lower_order(1).orders == (2, 2)
    with pytest.raises(ValueError):
        fs.lower_order(2)
    with pytest.raises(ValueError):
        FunctionSpace(BasisSpecs("legendre", 0)).lower_order(0)


@pytest.mark.parametrize("orders", [(3,), (2, 4), (2, 2, 3)])
def test_integration_space_is_the_tensor_product(orders):
    # reference tests/test_integration_space.py:16-116
    space = IntegrationSpace(*(IntegrationSpecs(o, "gauss") for o in orders))
    assert space.dimension == len(orders) and space.orders == orders
    nodes, weights = space.nodes(), space.weights()
    assert nodes.shape == (len(orders),) + tuple(o + 1 for o in orders) and weights.shape == nodes.shape[1:]
    ones = [s.nodes() for s in space.integration_specs]
    grids = np.meshgrid(*ones, indexing="ij")
    for k in range(len(orders)):
        assert np.array_equal(nodes[k], grids[k])
    w = np.ones(())
    for s in space.integration_specs:
        w = w[..., None] * s.weights()
    assert np.array_equal(weights, w)  # built in ascending dimension order like integration_objects.c:683-702
    # integrates a separable polynomial exactly
    f = np.prod([nodes[k] ** 2 for k in range(len(orders))], axis=0)
    assert np.sum(f * weights) == pytest.approx((2.0 / 3.0) ** len(orders), rel=1e-13)


@pytest.mark.parametrize("btype", ALL_TYPES)
@pytest.mark.parametrize("order", [1, 2, 5])
def test_basis_values_partition_and_derivative(btype, order):
    specs = BasisSpecs(btype, order)
    x = np.linspace(-1, 1, 17)
    v, d = specs.values(x), specs.derivatives(x)
    assert v.shape == d.shape == (17, order + 1)
    if btype != BasisType.LEGENDRE:
        assert np.allclose(v.sum(axis=1), 1.0, atol=1e-13) and np.allclose(d.sum(axis=1), 0.0, atol=1e-11)
    h = 1e-6
    fd = (specs.values(x + h) - specs.values(x - h)) / (2 * h)
    assert np.allclose(d, fd, atol=1e-6 * max(1.0, np.abs(d).max()))


@pytest.mark.parametrize("types", list(itertools.product(ALL_TYPES[:3], ALL_TYPES[3:])))
def test_evaluate_is_the_outer_product_of_1d_values(types):
    # reference tests/test_function_space.py:13-87 (exact equality in 1-D / 2-D)
    specs = [BasisSpecs(t, o) for t, o in zip(types, (3, 2))]
    fs = FunctionSpace(*specs)
    rng = np.random.default_rng(7)
    x, y = rng.uniform(-1, 1, (2, 4, 5))
    out = fs.evaluate(x, y)
    assert out.shape == (4, 5, 4, 3)
    assert np.array_equal(out, specs[0].values(x)[..., :, None] * specs[1].values(y)[..., None, :])
    buf = np.empty_like(out)
    assert fs.evaluate(x, y, out=buf) is buf and np.array_equal(buf, out)
    with pytest.raises(TypeError):
        fs.evaluate(x)
    with pytest.raises(ValueError):
        fs.evaluate(x, y[:2])


@pytest.mark.parametrize("orders", [(2,), (3, 1), (1, 2, 2)])
def test_values_at_integration_nodes_equals_evaluate_at_nodes(orders):
    # reference tests/test_function_space.py:90-107
    fs = FunctionSpace(*(BasisSpecs("lagrange-gauss-lobatto", o) for o in orders))
    space = IntegrationSpace(*(IntegrationSpecs(o + 2, "gauss") for o in orders))
    nodes = space.nodes()
    a = fs.values_at_integration_nodes(space)
    b = fs.evaluate(*[nodes[k] for k in range(len(orders))])
    assert a.shape == b.shape and np.allclose(a, b, rtol=0, atol=1e-13)
    with pytest.raises(ValueError):
        fs.values_at_integration_nodes(IntegrationSpace(IntegrationSpecs(2), IntegrationSpecs(2), IntegrationSpecs(2), IntegrationSpecs(2)))


def test_registries_record_usage_and_clear():
    ireg, breg = IntegrationRegistry(), BasisRegistry()
    assert ireg.usage() == () and breg.usage() == ()
    fs = FunctionSpace(BasisSpecs("legendre", 2), BasisSpecs("bernstein", 3))
    space = IntegrationSpace(IntegrationSpecs(4), IntegrationSpecs(5, "gauss-lobatto"))
    fs.values_at_integration_nodes(space, integration_registry=ireg, basis_registry=breg)
    assert set(ireg.usage()) == set(space.integration_specs)
    assert set(breg.usage()) == set(zip(fs.basis_specs, space.integration_specs))
    assert IntegrationSpecs(4).nodes(ireg) is IntegrationSpecs(4).nodes(ireg)  # cached object, like the C registry's pointer
    ireg.clear()
    breg.clear()
    assert ireg.usage() == () and breg.usage() == ()
    with pytest.raises(TypeError):
        IntegrationSpecs(2).nodes(breg)


def test_dofs_container_and_reconstruct_at_points():
    # reference tests/test_dofs.py:19-39
    fs = FunctionSpace(BasisSpecs("legendre", 2), BasisSpecs("legendre", 4), BasisSpecs("legendre", 5))
    dofs = DegreesOfFreedom(fs)
    assert dofs.function_space == fs and dofs.n_dofs == 90 and dofs.shape == (3, 5, 6) and not dofs.values.any()
    vals = np.random.default_rng(1234).random((3, 5, 6))
    dofs.values = vals
    assert np.array_equal(dofs.values, vals)
    dofs.values = np.zeros(90)
    assert not dofs.values.any()
    with pytest.raises(ValueError):
        dofs.values = np.zeros(89)
    # host-side reconstruction of a polynomial that lies in the space
    fs1 = FunctionSpace(BasisSpecs("lagrange-uniform", 3))
    xs = np.linspace(-1, 1, 4)
    d1 = DegreesOfFreedom(fs1, xs**3 - xs)
    pts = np.linspace(-1, 1, 11)
    assert np.allclose(reconstruct(d1, pts), pts**3 - pts, atol=1e-14)


def test_geometry_authoring_formats():
    q = Quad.from_corners((0, 0), (2, 0), (2.5, 1.5), (0, 1))
    assert q.ndim_physical == 2 and q.ndim_reference == 2 and q.function_space.orders == (1, 1)
    g = q.geometry_dofs.reshape(2, 2, 2)  # [c][i0][i1], corner at (2*i0-1, 2*i1-1)
    assert g[:, 0, 0].tolist() == [0, 0] and g[:, 1, 0].tolist() == [2, 0] and g[:, 1, 1].tolist() == [2.5, 1.5] and g[:, 0, 1].tolist() == [0, 1]
    x, y = q.sample(np.array([-1.0, 1.0, 0.0]), np.array([-1.0, 1.0, 0.0]))
    assert np.allclose(x, [0, 2.5, 1.125]) and np.allclose(y, [0, 1.5, 0.625])
    ln = Line((0.0, 1.0), (2.0, 3.0))
    assert ln.function_space.basis_specs[0].type == "bernstein" and ln.geometry_dofs.tolist() == [[0, 2], [1, 3]]
    with pytest.raises(ValueError):
        Line((0.0, 1.0))


def test_numeric_calls_have_no_cpu_fallback():
    import interplib_b200 as ib

    if ib.device_count() > 0:
        pytest.skip("a GPU is visible here")
    fs = FunctionSpace(BasisSpecs("legendre", 2))
    with pytest.raises(ib.InterpError):
        compute_mass_matrix(fs, fs, IntegrationSpace(IntegrationSpecs(3)))
