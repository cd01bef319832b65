"""Drop-in mirror of the reference's per-element API for the hot path (``interplib`` / ``interplib._interp``).

Same names, argument meaning, array shapes and exception types as the reference's CPython extension
(``/root/reference/src/python/*.c``, stub ``python/interplib/_interp.pyi``) and its NumPy glue
(``python/interplib/integration.py``, ``degrees_of_freedom.py``); every method cites what it mirrors.
Nothing is computed in Python: table-only methods call the host C layer of ``libinterp_b200.so``, every
numeric method is a batch of ONE element through the batched C ABI on the GPU (``interplib_b200.batch``).
There is no CPU implementation behind the numeric calls -- without a B200 they raise.

The element-batched counterparts (the performance path) are ``interplib_b200.batch.Context`` methods; the
objects below expose ``.spec`` tuples so the two layers share descriptions.
"""
from __future__ import annotations

from enum import StrEnum
from typing import Callable, Sequence

import numpy as np

from . import batch
from ._lib import BASIS_TYPES, RULE_TYPES

__all__ = [
    "BasisType", "IntegrationMethod", "IntegrationRegistry", "IntegrationSpecs", "IntegrationSpace", "BasisRegistry",
    "BasisSpecs", "FunctionSpace", "DegreesOfFreedom", "CoordinateMap", "SpaceMap", "compute_mass_matrix",
    "compute_gradient_mass_matrix", "projection_l2_dual", "projection_l2_primal", "integrate_callable", "reconstruct", "reconstruct_batch", "DegreesOfFreedomBatch",
    "projection_l2_primal_batch", "incidence_matrix",
    "compute_dual_degrees_of_freedom", "DEFAULT_INTEGRATION_REGISTRY", "DEFAULT_BASIS_REGISTRY",
]


class BasisType(StrEnum):
    """python/interplib/enum_type.py:7-15 (member spellings kept, including the reference's typos)."""

    LEGENDRE = "legendre"
    BERNSTEIN = "bernstein"
    LAGRANGE_UNIFORM = "lagrange-uniform"
    LAGRNAGE_GAUSS = "lagrange-gauss"
    LAGRNAGE_GAUSS_LOBATTO = "lagrange-gauss-lobatto"
    LAGRANGE_CHEBYSHEV_GAUSS = "lagrange-chebyshev-gauss"


class IntegrationMethod(StrEnum):
    """python/interplib/enum_type.py:31-35"""

    GAUSS = "gauss"
    GAUSS_LOBATTO = "gauss-lobatto"


def _ctx() -> batch.Context:
    return batch.default_context()


# ------------------------------------------------------------------------------------------------ registries
class IntegrationRegistry:
    """Cache of 1-D rules (src/python/integration_objects.c:30-181; C cache src/integration/integration_rules.c:146-341).
    Host copies of the nodes/weights live here; the device copies live in the GPU context's table cache."""

    def __init__(self, *args, **kwargs):
        if args or kwargs:
            raise TypeError("IntegrationRegistry takes no arguments.")
        self._rules: dict[tuple[str, int], tuple[np.ndarray, np.ndarray]] = {}

    def _get(self, method: str, order: int):
        key = (method, order)
        hit = self._rules.get(key)
        if hit is None:
            x, w, _ = batch.rule_1d(method, order)
            x.setflags(write=False)
            w.setflags(write=False)
            hit = self._rules[key] = (x, w)
        return hit

    def usage(self) -> tuple["IntegrationSpecs", ...]:
        return tuple(IntegrationSpecs(o, m) for m, o in self._rules)

    def clear(self) -> None:
        self._rules.clear()


class BasisRegistry:
    """Cache of 1-D basis tables keyed by (basis specs, rule specs) (src/python/basis_objects.c:75-240;
    src/basis/basis_set.c:212-287)."""

    def __init__(self, *args, **kwargs):
        if args or kwargs:
            raise TypeError("BasisRegistry takes no arguments.")
        self._sets: dict[tuple, tuple[np.ndarray, np.ndarray]] = {}

    def _get(self, basis: tuple[str, int], rule: tuple[str, int]):
        key = (basis, rule)
        hit = self._sets.get(key)
        if hit is None:
            hit = self._sets[key] = batch.basis_table_1d(basis, rule)
        return hit

    def usage(self) -> tuple[tuple["BasisSpecs", "IntegrationSpecs"], ...]:
        return tuple((BasisSpecs(bt, bo), IntegrationSpecs(ro, rm)) for (bt, bo), (rm, ro) in self._sets)

    def clear(self) -> None:
        self._sets.clear()


DEFAULT_INTEGRATION_REGISTRY = IntegrationRegistry()
DEFAULT_BASIS_REGISTRY = BasisRegistry()


def _check_registry(reg, cls):
    if not isinstance(reg, cls):
        raise TypeError(f"Expected {cls.__name__}, got {type(reg).__name__}")
    return reg


# ------------------------------------------------------------------------------------------------ specs
class IntegrationSpecs:
    """src/python/integration_objects.c:213-510: ``IntegrationSpecs(order, /, method="gauss")``."""

    __slots__ = ("_order", "_method")

    def __init__(self, order: int, /, method: str = "gauss"):
        method = str(method)
        if method not in RULE_TYPES:
            raise ValueError(f'Unknown integration method "{method}".')
        if not isinstance(order, (int, np.integer)) or isinstance(order, bool):
            raise TypeError("Order must be an integer.")
        if order < 0:
            raise ValueError("Order must be a non-negative integer.")
        self._order, self._method = int(order), method

    order = property(lambda self: self._order)
    method = property(lambda self: self._method)

    @property
    def accuracy(self) -> int:
        """Highest polynomial order integrated exactly (src/integration/integration_rules.c:64-71)."""
        return batch.rule_1d(self._method, self._order)[2]

    @property
    def spec(self) -> tuple[str, int]:
        return (self._method, self._order)

    def nodes(self, registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY) -> np.ndarray:
        return _check_registry(registry, IntegrationRegistry)._get(self._method, self._order)[0]

    def weights(self, registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY) -> np.ndarray:
        return _check_registry(registry, IntegrationRegistry)._get(self._method, self._order)[1]

    def __eq__(self, other):
        return isinstance(other, IntegrationSpecs) and self.spec == other.spec

    def __hash__(self):
        return hash(self.spec)

    def __repr__(self):
        return f"IntegrationSpecs({self._order}, {self._method!r})"


class BasisSpecs:
    """src/python/basis_objects.c:243-553: ``BasisSpecs(basis_type, order, /)``."""

    __slots__ = ("_type", "_order")

    def __init__(self, basis_type: str, order: int, /):
        basis_type = str(basis_type)
        if basis_type not in BASIS_TYPES:
            raise ValueError(f"Unknown basis type string: {basis_type}")
        if order < 0:
            raise ValueError(f"Order must be positive, but was given as {order}.")
        self._type, self._order = basis_type, int(order)

    order = property(lambda self: self._order)
    type = property(lambda self: self._type)  # C name (basis_objects.c:296)
    basis_type = property(lambda self: self._type)  # name in the typing stub (_interp.pyi:495)

    @property
    def spec(self) -> tuple[str, int]:
        return (self._type, self._order)

    def values(self, x, /) -> np.ndarray:
        """Basis values at arbitrary points -> x.shape + (order+1,) (basis_objects.c:351-428, src/basis/basis_set.c:462-489)."""
        return batch.basis_at_points(self.spec, np.asarray(x, dtype=np.float64))

    def derivatives(self, x, /) -> np.ndarray:
        """basis_objects.c:430-491, src/basis/basis_set.c:491-556"""
        return batch.basis_at_points(self.spec, np.asarray(x, dtype=np.float64), derivative=True)

    def __eq__(self, other):
        return isinstance(other, BasisSpecs) and self.spec == other.spec

    def __hash__(self):
        return hash(self.spec)

    def __repr__(self):
        return f"BasisSpecs({self._type!r}, {self._order})"


# ------------------------------------------------------------------------------------------------ spaces
class IntegrationSpace:
    """src/python/integration_objects.c:512-876: tensor product of 1-D rules."""

    def __init__(self, *integration_specs: IntegrationSpecs, **kwargs):
        if kwargs:
            raise TypeError("Constructor takes no keyword arguments.")
        if not integration_specs:
            raise TypeError("Expected at least one argument.")
        for s in integration_specs:
            if not isinstance(s, IntegrationSpecs):
                raise TypeError(f"Expected IntegrationSpecs, got {type(s).__name__}")
        self._specs = tuple(integration_specs)

    dimension = property(lambda self: len(self._specs))
    integration_specs = property(lambda self: self._specs)
    orders = property(lambda self: tuple(s.order for s in self._specs))

    @property
    def rules(self) -> list[tuple[str, int]]:
        """``[(method, order), ...]`` -- the description the batched API takes."""
        return [s.spec for s in self._specs]

    @property
    def shape(self) -> tuple[int, ...]:
        return tuple(s.order + 1 for s in self._specs)

    def nodes(self, registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY, /) -> np.ndarray:
        """[d, m0, m1, ...] (integration_objects.c:731-814)"""
        _check_registry(registry, IntegrationRegistry)
        for s in self._specs:
            registry._get(*s.spec)
        return batch.integration_space_nodes(self.rules)

    def weights(self, registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY, /) -> np.ndarray:
        """[m0, m1, ...], product built in ascending dimension order (integration_objects.c:633-716)"""
        _check_registry(registry, IntegrationRegistry)
        for s in self._specs:
            registry._get(*s.spec)
        return batch.integration_space_weights(self.rules)

    def __eq__(self, other):
        return isinstance(other, IntegrationSpace) and self._specs == other._specs

    def __hash__(self):
        return hash(self._specs)

    def __repr__(self):
        return "IntegrationSpace(" + ", ".join(map(repr, self._specs)) + ")"


class FunctionSpace:
    """src/python/function_space_objects.c:16-636: tensor product of 1-D bases."""

    def __init__(self, *basis_specs: BasisSpecs, **kwargs):
        if kwargs:
            raise TypeError("Constructor takes no keyword arguments.")
        if not basis_specs:
            raise TypeError("Constructor requires at least one argument.")
        for i, s in enumerate(basis_specs):
            if not isinstance(s, BasisSpecs):
                raise TypeError(f"Argument {i} was not a BasisSpec, but was instead {s!r}.")
        self._specs = tuple(basis_specs)

    dimension = property(lambda self: len(self._specs))
    basis_specs = property(lambda self: self._specs)
    orders = property(lambda self: tuple(s.order for s in self._specs))

    @property
    def basis(self) -> list[tuple[str, int]]:
        """``[(type, order), ...]`` -- the description the batched API takes."""
        return [s.spec for s in self._specs]

    @property
    def shape(self) -> tuple[int, ...]:
        return tuple(s.order + 1 for s in self._specs)

    @property
    def n_basis(self) -> int:
        return int(np.prod(self.shape))

    def evaluate(self, *x, out: np.ndarray | None = None) -> np.ndarray:
        """Tensor basis at arbitrary points -> x.shape + (n0, n1, ...) (function_space_objects.c:142-276,
        src/basis/basis_set.c:577-633)."""
        if len(x) != self.dimension:
            raise TypeError("evaluate() takes exactly the same number of positional arguments as the dimension count "
                            f"({self.dimension}), but {len(x)} were given.")
        xs = [np.ascontiguousarray(v, dtype=np.float64) for v in x]
        for v in xs[1:]:
            if v.shape != xs[0].shape:
                raise ValueError("All input coordinate arrays must have the same shape.")
        res = batch.function_space_evaluate(self.basis, xs)
        if out is not None:
            if out.shape != res.shape or out.dtype != np.float64 or not out.flags.c_contiguous:
                raise ValueError("Output array must be a C-contiguous float64 array of shape " + str(res.shape))
            out[...] = res
            return out
        return res

    def values_at_integration_nodes(self, integration: IntegrationSpace, /, *,
                                    integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                                    basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY) -> np.ndarray:
        """[m0, .., n0, ..] (function_space_objects.c:301-474)"""
        if not isinstance(integration, IntegrationSpace):
            raise TypeError(f"Expected IntegrationSpace, got {type(integration).__name__}")
        if integration.dimension != self.dimension:
            raise ValueError("Function space and integration space must have the same number of dimensions (function space "
                             f"had {self.dimension}, but integration space had {integration.dimension}).")
        _touch(self, integration, integration_registry, basis_registry)
        return batch.values_at_integration_nodes(self.basis, integration.rules)

    def lower_order(self, idim: int) -> "FunctionSpace":
        """function_space_objects.c:490-530"""
        if idim < 0 or idim >= self.dimension:
            raise ValueError(f"Dimension {idim} is out of bounds for a function space with {self.dimension} dimensions.")
        if self._specs[idim].order == 0:
            raise ValueError(f"Cannot lower order of dimension {idim} as it has order 0.")
        specs = list(self._specs)
        specs[idim] = BasisSpecs(specs[idim].type, specs[idim].order - 1)
        return FunctionSpace(*specs)

    def __eq__(self, other):
        return isinstance(other, FunctionSpace) and self._specs == other._specs

    def __hash__(self):
        return hash(self._specs)

    def __repr__(self):
        return "FunctionSpace(" + ", ".join(map(repr, self._specs)) + ")"


def _touch(fs: FunctionSpace, ispace: IntegrationSpace, ireg: IntegrationRegistry, breg: BasisRegistry) -> None:
    """Record the (basis, rule) pairs a call uses, as the reference's registries would (usage() parity)."""
    _check_registry(ireg, IntegrationRegistry)
    _check_registry(breg, BasisRegistry)
    for b, r in zip(fs.basis_specs, ispace.integration_specs):
        ireg._get(*r.spec)
        breg._get(b.spec, r.spec)


# ------------------------------------------------------------------------------------------------ DoFs
class DegreesOfFreedom:
    """src/python/degrees_of_freedom.c:33-1015: ``DegreesOfFreedom(function_space, values=None, /)``."""

    def __init__(self, function_space: FunctionSpace, values=None, /):
        if not isinstance(function_space, FunctionSpace):
            raise TypeError(f"Expected FunctionSpace, got {type(function_space).__name__}")
        self._fs = function_space
        self._values = np.zeros(function_space.shape)
        if values is not None:
            self.values = values

    function_space = property(lambda self: self._fs)
    n_dofs = property(lambda self: self._fs.n_basis)
    shape = property(lambda self: self._fs.shape)

    @property
    def values(self) -> np.ndarray:
        return self._values

    @values.setter
    def values(self, value) -> None:
        arr = np.asarray(value, dtype=np.float64)
        if arr.shape != self.shape and arr.shape != (self.n_dofs,):
            raise ValueError(f"Values must either be flat with {self.n_dofs} elements or have exact correct "
                             f"{len(self.shape)}-dimensional shape.")
        self._values[...] = arr.reshape(self.shape)

    def _check_space(self, integration_space: IntegrationSpace, out):
        if not isinstance(integration_space, IntegrationSpace):
            raise TypeError(f"Expected IntegrationSpace, got {type(integration_space).__name__}")
        ndim = self._fs.dimension
        if integration_space.dimension != ndim:
            raise ValueError(f"Expected integration space with {ndim} dimensions, but it had only {integration_space.dimension}.")
        if out is not None:
            if not isinstance(out, np.ndarray) or out.dtype != np.float64 or not out.flags.c_contiguous:
                raise ValueError("Output array must be a C-contiguous array of doubles.")
            if out.ndim != ndim:
                raise ValueError(f"Output array must have {ndim} dimensions, but it had {out.ndim}.")
            for k, (m, a) in enumerate(zip(integration_space.shape, out.shape)):
                if m != a:
                    raise ValueError("Output array must have the exact same shape as the integration space, but dimension "
                                     f"{k} did not match (integration space: {m}, array: {a}).")

    def _reconstruct(self, integration_space, mask, ireg, breg, out):
        self._check_space(integration_space, out)
        _touch(self._fs, integration_space, ireg, breg)
        res = _ctx().reconstruct(self._fs.basis, integration_space.rules, self._values.reshape(1, -1), [mask])
        res = res.reshape(integration_space.shape)
        if out is not None:
            out[...] = res
            return out
        return res

    def reconstruct_at_integration_points(self, integration_space: IntegrationSpace,
                                          integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                                          basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY, *, out=None) -> np.ndarray:
        """[m0, m1, ...] (degrees_of_freedom.c:426-510 -> compute_integration_point_values :232-262): K1, batch of one."""
        return self._reconstruct(integration_space, 0, integration_registry, basis_registry, out)

    def reconstruct_derivative_at_integration_points(self, integration_space: IntegrationSpace, idim: Sequence[int],
                                                     integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                                                     basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY, *, out=None) -> np.ndarray:
        """Derivative along every reference dimension listed in ``idim`` (each at most once)
        (degrees_of_freedom.c:622-722, index parsing :512-592)."""
        ndim = self._fs.dimension
        if isinstance(idim, (int, np.integer)):
            idim = [int(idim)]
        mask = 0
        for i, idx in enumerate(idim):
            if not isinstance(idx, (int, np.integer)):
                raise TypeError(f"Expected a sequence of integers, but got a non-integer at index {i}.")
            if idx < 0 or idx >= ndim:
                raise ValueError(f"Expected an index between 0 and {ndim - 1}, but got {idx}.")
            if mask & (1 << idx):
                raise ValueError(f"Expected each index to appear only once, but got it twice at index {idx}.")
            mask |= 1 << int(idx)
        return self._reconstruct(integration_space, mask, integration_registry, basis_registry, out)

    def derivative(self, idim: int, /) -> "DegreesOfFreedom":
        """DoFs of the derivative along reference dimension ``idim`` in the space one order lower along it
        (degrees_of_freedom.c:853-945; ib200_dof_derivative_batch with a batch of one, bit-identical)."""
        ndim = self._fs.dimension
        if not isinstance(idim, (int, np.integer)):
            raise TypeError(f"Expected an integer index, got {type(idim).__name__}.")
        if idim < 0 or idim >= ndim:
            raise ValueError(f"Expected an index between 0 and {ndim - 1}, but got {idim}.")
        if self._fs.basis[idim][1] == 0:
            raise ValueError(f"Cannot compute the derivative of a function with order 0 in dimension {idim}.")
        lower = self._fs.lower_order(int(idim))
        vals = _ctx().dof_derivative(self._fs.basis, self._values.reshape(1, -1), int(idim))
        return DegreesOfFreedom(lower, vals.reshape(lower.shape))


def incidence_matrix(basis_specs: "BasisSpecs", /) -> np.ndarray:
    """``[n, n+1]`` matrix taking the DoFs of a 1-D function to those of its derivative, one order lower
    (src/python/incidence.c:6-110; host table layer, bit-identical)."""
    if not isinstance(basis_specs, BasisSpecs):
        raise TypeError(f"Expected BasisSpecs, got {type(basis_specs).__name__}")
    if basis_specs.order == 0:
        raise ValueError("Cannot compute the incidence matrix for a zero-dimensional basis.")
    return batch.incidence_matrix(basis_specs.spec)


# ------------------------------------------------------------------------------------------------ geometry
class CoordinateMap:
    """One physical coordinate at the integration points: values + d reference gradients, storage [1+d][Q]
    (src/python/mappings.c:8-280, mappings.h:6-12)."""

    def __init__(self, dofs: DegreesOfFreedom, integration_space: IntegrationSpace,
                 integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                 basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY, /):
        if not isinstance(dofs, DegreesOfFreedom):
            raise TypeError(f"Expected DegreesOfFreedom, got {type(dofs).__name__}")
        dofs._check_space(integration_space, None)
        _touch(dofs.function_space, integration_space, integration_registry, basis_registry)
        self._ispace = integration_space
        self._dofs = dofs
        d = integration_space.dimension
        masks = [0] + [1 << k for k in range(d)]
        res = _ctx().reconstruct(dofs.function_space.basis, integration_space.rules, dofs.values.reshape(1, -1), masks)
        self._data = res.reshape((1 + d,) + integration_space.shape)
        self._data.setflags(write=False)

    dimension = property(lambda self: self._ispace.dimension)
    integration_space = property(lambda self: self._ispace)

    @property
    def values(self) -> np.ndarray:
        return self._data[0]

    def gradient(self, idim: int, /) -> np.ndarray:
        if idim < 0 or idim >= self.dimension:
            raise ValueError(f"Expected dimension index in range [0, {self.dimension}), but got {idim}.")
        return self._data[1 + idim]


class SpaceMap:
    """Reference -> physical map at the integration points: Jacobian, Givens QR, det = prod diag(R), left inverse
    (src/python/mappings.c:304-787, src/operations/matrices.c:15-141): K0, batch of one."""

    def __init__(self, *coordinates: CoordinateMap, **kwargs):
        if kwargs:
            raise TypeError("SpaceMap takes no keyword arguments.")
        if not coordinates:
            raise TypeError("SpaceMap requires at least one argument.")
        for c in coordinates:
            if not isinstance(c, CoordinateMap):
                raise TypeError(f"Expected a CoordinateMap, but got a {type(c).__name__}.")
        first = coordinates[0]
        ndim, n_phys = first.dimension, len(coordinates)
        if n_phys < ndim:
            raise ValueError(f"Can not construct a space map from reference domain with {ndim} dimensions to {n_phys} physical "
                             "dimensions. The number of physical dimensions must be equal to, or greater than the number of "
                             "reference dimensions.")
        for i, c in enumerate(coordinates[1:], 1):
            if c.dimension != ndim:
                raise ValueError("Expected all coordinate maps to have the same number of dimensions, but "
                                 f"got {ndim} and {c.dimension}.")
            if c.integration_space != first.integration_space:
                raise ValueError("Expected all coordinate maps to have the same integration space, but the first and "
                                 f"{i} space have different ones.")
            if c._dofs.function_space != first._dofs.function_space:
                # the batched geometry kernel takes one geometry space for all coordinates (the reference's Line/Quad
                # domains build them that way, python/interplib/domains.py:34-50)
                raise ValueError("Expected all coordinate maps to use the same function space.")
        self._coords = tuple(coordinates)
        self._ispace = first.integration_space
        geom = np.stack([c._dofs.values.reshape(-1) for c in coordinates])[None]
        _, det, jinv = _ctx().geometry(first._dofs.function_space.basis, self._ispace.rules, geom, n_phys, want=("det", "jinv"))
        self._det = det.reshape(self._ispace.shape)
        self._inv = jinv.reshape(self._ispace.shape + (ndim, n_phys))
        self._det.setflags(write=False)
        self._inv.setflags(write=False)

    integration_space = property(lambda self: self._ispace)
    input_dimensions = property(lambda self: self._ispace.dimension)
    output_dimensions = property(lambda self: len(self._coords))
    determinant = property(lambda self: self._det)
    inverse_map = property(lambda self: self._inv)  # C name (mappings.c:769)
    inverse_transform = property(lambda self: self._inv)  # name in the typing stub (_interp.pyi:905)

    def coordinate_map(self, idx: int) -> CoordinateMap:
        if idx < 0 or idx >= len(self._coords):
            raise ValueError(f"Expected dimension index in range [0, {len(self._coords)}), but got {idx}.")
        return self._coords[idx]

    @property
    def geometry_dofs(self) -> np.ndarray:
        """[n_phys][N_g] coordinate DoFs (what the fused batched kernels take instead of a SpaceMap)."""
        return np.stack([c._dofs.values.reshape(-1) for c in self._coords])


# ------------------------------------------------------------------------------------------------ mass matrices
def _integration_args(integration):
    if isinstance(integration, IntegrationSpace):
        return integration, None
    if isinstance(integration, SpaceMap):
        return integration.integration_space, integration
    raise TypeError(f"Integration space or space map must be passed, instead {type(integration).__name__} object was passed.")


def _check_spaces(space_in, space_out, ispace):
    for s in (space_in, space_out):
        if not isinstance(s, FunctionSpace):
            raise TypeError(f"Expected FunctionSpace, got {type(s).__name__}")
    if space_out.dimension != space_in.dimension or ispace.dimension != space_in.dimension:
        raise ValueError("Function spaces must have the same dimensionality (space in: "
                         f"{space_in.dimension}, space out: {space_out.dimension}, integration space: {ispace.dimension}).")


def compute_mass_matrix(space_in: FunctionSpace, space_out: FunctionSpace, integration, /, *,
                        integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                        basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY) -> np.ndarray:
    """``[N_out, N_in]`` mass matrix on the reference domain (IntegrationSpace) or on the mapped element (SpaceMap)
    (src/python/mass_matrices.c:161-339): K2, batch of one."""
    ispace, smap = _integration_args(integration)
    _check_spaces(space_in, space_out, ispace)
    _touch(space_in, ispace, integration_registry, basis_registry)
    _touch(space_out, ispace, integration_registry, basis_registry)
    weight = None if smap is None else smap.determinant.reshape(1, -1)
    return _ctx().mass_matrix(space_in.basis, space_out.basis, ispace.rules, 1, weight=weight)[0]


def compute_gradient_mass_matrix(space_in: FunctionSpace, space_out: FunctionSpace, integration, /, idim_in: int, idim_out: int, *,
                                 integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                                 basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY) -> np.ndarray:
    """Mass matrix with the input basis differentiated along reference dimension ``idim_in`` and weighted by
    ``det * Jinv[idim_in][idim_out]`` (src/python/mass_matrices.c:341-565)."""
    ispace, smap = _integration_args(integration)
    n_coords = ispace.dimension if smap is None else smap.output_dimensions
    if idim_in < 0 or idim_in >= ispace.dimension:
        raise ValueError(f"Index {idim_in} out of bounds for input space with {ispace.dimension} dimensions.")
    if idim_out < 0 or idim_out >= n_coords:
        raise ValueError(f"Index {idim_out} out of bounds for output space with {n_coords} dimensions.")
    _check_spaces(space_in, space_out, ispace)
    if smap is None and idim_in != idim_out:
        return np.zeros((space_out.n_basis, space_in.n_basis))  # mass_matrices.c:463-474
    _touch(space_in, ispace, integration_registry, basis_registry)
    _touch(space_out, ispace, integration_registry, basis_registry)
    weight = None
    if smap is not None:
        weight = (smap.determinant * smap.inverse_map[..., idim_in, idim_out]).reshape(1, -1)  # mass_matrices.c:516-518
    return _ctx().mass_matrix(space_in.basis, space_out.basis, ispace.rules, 1, weight=weight, deriv_dim_in=idim_in)[0]


# ------------------------------------------------------------------------------------------------ NumPy-glue layer
def _prepare_integration(integration, registry):
    """python/interplib/integration.py:44-67"""
    if isinstance(integration, IntegrationSpace):
        return integration.nodes(registry), None
    if isinstance(integration, SpaceMap):
        nodes = np.array([integration.coordinate_map(i).values for i in range(integration.output_dimensions)])
        return nodes, integration.determinant
    raise TypeError(f"Only {IntegrationSpace} or {SpaceMap} can be passed, but instead {type(integration)} was passed.")


def integrate_callable(func: Callable, integration, /, *, registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY) -> float:
    """python/interplib/integration.py:70-103"""
    nodes, det = _prepare_integration(integration, registry)
    ispace, _ = _integration_args(integration)
    weights = ispace.weights(registry)
    if det is not None:
        weights = weights * det
    return float(np.sum(np.asarray(func(*[nodes[i, ...] for i in range(nodes.shape[0])])) * weights))


def projection_l2_dual(func: Callable, function_space: FunctionSpace, integration, /, *,
                       integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                       basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY) -> DegreesOfFreedom:
    """Dual DoFs ``b_i = sum_q f(x_q) W_q det_q phi_i(q)`` (python/interplib/integration.py:106-168): the callable is
    sampled on the host at the (mapped) nodes, the contraction is K3 on the GPU, batch of one."""
    nodes, det = _prepare_integration(integration, integration_registry)
    ispace, _ = _integration_args(integration)
    if function_space.dimension != ispace.dimension:
        raise ValueError("Function space and integration space must have the same number of dimensions (function space "
                         f"had {function_space.dimension}, but integration space had {ispace.dimension}).")
    fvals = np.broadcast_to(np.asarray(func(*[nodes[i, ...] for i in range(nodes.shape[0])]), dtype=np.float64), ispace.shape)
    _touch(function_space, ispace, integration_registry, basis_registry)
    dual = _ctx().l2_dual(function_space.basis, ispace.rules, np.ascontiguousarray(fvals).reshape(1, -1),
                          det=None if det is None else det.reshape(1, -1))
    return DegreesOfFreedom(function_space, dual[0])


def projection_l2_primal(func: Callable, function_space: FunctionSpace, integration, /, *,
                         integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                         basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY) -> DegreesOfFreedom:
    """Primal DoFs ``M u = b`` (python/interplib/integration.py:171-223; the reference calls np.linalg.solve, here the
    batched Cholesky kernel)."""
    dual = projection_l2_dual(func, function_space, integration, integration_registry=integration_registry,
                              basis_registry=basis_registry)
    mass = compute_mass_matrix(function_space, function_space, integration, integration_registry=integration_registry,
                               basis_registry=basis_registry)
    primal = _ctx().spd_solve(mass[None], dual.values.reshape(1, -1))
    return DegreesOfFreedom(function_space, primal[0])


def reconstruct(dof: DegreesOfFreedom, *x) -> np.ndarray:
    """Function values at arbitrary points (python/interplib/degrees_of_freedom.py:10-37)."""
    xs = [np.ascontiguousarray(v, dtype=np.float64) for v in x]
    for v in xs[1:]:
        if v.shape != xs[0].shape:
            raise ValueError("All input coordinate arrays must have the same shape.")
    phi = dof.function_space.evaluate(*xs).reshape((-1, dof.n_dofs))
    return np.sum(phi * dof.values.flatten(), axis=-1).reshape(xs[0].shape)


class DegreesOfFreedomBatch:
    """E stacked :class:`DegreesOfFreedom` of one function space, resident on the host (NumPy ``[E, n0, n1, ..]``) or on the
    device (``batch.DeviceArray``).  The element-batched counterpart of the reference's per-element object: its methods are
    one GPU launch for all E elements, and on a device-resident batch inputs and results stay in HBM -- chains such as
    ``projection_l2_primal_batch(...).reconstruct_at_integration_points(...)`` never cross PCIe until ``to_host()``."""

    def __init__(self, function_space: FunctionSpace, values):
        if not isinstance(function_space, FunctionSpace):
            raise TypeError(f"Expected FunctionSpace, got {type(function_space).__name__}")
        self._fs = function_space
        if isinstance(values, batch.DeviceArray):
            self._values = values.reshape((-1,) + function_space.shape)
        else:
            arr = np.ascontiguousarray(values, dtype=np.float64)
            if arr.size % function_space.n_basis:
                raise ValueError(f"values must hold a multiple of {function_space.n_basis} entries")
            self._values = arr.reshape((-1,) + function_space.shape)

    function_space = property(lambda self: self._fs)
    n_elements = property(lambda self: self._values.shape[0])
    on_device = property(lambda self: isinstance(self._values, batch.DeviceArray))
    values = property(lambda self: self._values)

    def to_device(self) -> "DegreesOfFreedomBatch":
        return self if self.on_device else DegreesOfFreedomBatch(self._fs, _ctx().to_device(self._values))

    def to_host(self) -> "DegreesOfFreedomBatch":
        return DegreesOfFreedomBatch(self._fs, self._values.to_host()) if self.on_device else self

    def __getitem__(self, e: int) -> DegreesOfFreedom:
        """Element e as the reference's per-element object (host batches)."""
        if self.on_device:
            raise TypeError("index a host batch (to_host() first)")
        return DegreesOfFreedom(self._fs, self._values[e])

    def reconstruct_at_integration_points(self, integration_space: IntegrationSpace, idim: Sequence[Sequence[int]] = ((),)):
        """``[E, n_fields, m0, m1, ..]``: field k is the reconstruction differentiated once along every dimension listed in
        ``idim[k]`` (``()`` = the values) -- DegreesOfFreedom.reconstruct[_derivative]_at_integration_points for all
        elements in one launch (src/python/degrees_of_freedom.c:426-722)."""
        masks = []
        for dims in idim:
            m = 0
            for k in dims:
                if not 0 <= int(k) < self._fs.dimension or (m >> int(k)) & 1:
                    raise ValueError(f"Invalid or repeated derivative dimension {k}.")
                m |= 1 << int(k)
            masks.append(m)
        flat = self._values.reshape(self.n_elements, -1)
        out = _ctx().reconstruct(self._fs.basis, integration_space.rules, flat, masks)
        return out.reshape((self.n_elements, len(masks)) + tuple(o + 1 for o in integration_space.orders))

    def derivative(self, idim: int) -> "DegreesOfFreedomBatch":
        """DegreesOfFreedom.derivative for every element (src/python/degrees_of_freedom.c:853-945), bit-identical."""
        out = _ctx().dof_derivative(self._fs.basis, self._values.reshape(self.n_elements, -1), int(idim))
        return DegreesOfFreedomBatch(self._fs.lower_order(int(idim)), out)


def projection_l2_primal_batch(fvals, function_space: FunctionSpace, integration_space: IntegrationSpace, geom_dofs,
                               geom_space: FunctionSpace, n_phys: int | None = None) -> DegreesOfFreedomBatch:
    """projection_l2_primal (python/interplib/integration.py:171-223) for E elements at once: ``fvals[E, m0, m1, ..]`` are the
    samples of f at the mapped rule nodes, ``geom_dofs[E, n_phys, ...]`` the coordinate DoFs of ``geom_space`` (for
    ``Quad.from_corners`` elements: order-1 ``lagrange-uniform``).  NumPy inputs give a host batch, ``DeviceArray`` inputs a
    device-resident one (ib200_l2_project_mapped_batch: fast diagonalisation on bilinear quads, no mass matrix)."""
    n_phys = function_space.dimension if n_phys is None else int(n_phys)
    Q = int(np.prod([o + 1 for o in integration_space.orders]))
    dev = isinstance(fvals, batch.DeviceArray) or isinstance(geom_dofs, batch.DeviceArray)
    if dev:
        f2 = fvals.reshape(-1, Q)
        g2 = geom_dofs.reshape(f2.shape[0], n_phys, geom_space.n_basis)
    else:
        f2 = np.ascontiguousarray(fvals, dtype=np.float64).reshape(-1, Q)
        g2 = np.ascontiguousarray(geom_dofs, dtype=np.float64).reshape(f2.shape[0], n_phys, geom_space.n_basis)
    u = _ctx().l2_project_mapped(function_space.basis, integration_space.rules, geom_space.basis, g2, n_phys, f2)
    return DegreesOfFreedomBatch(function_space, u)


def reconstruct_batch(function_space: FunctionSpace, dof_values, *x, idim: Sequence[int] = ()) -> np.ndarray:
    """Element-batched :func:`reconstruct` on the GPU (ib200_reconstruct_points_batch):
    ``dof_values[E, n0, n1, ..]`` and one coordinate array ``x_k[E, ...]`` per dimension (same shape, leading axis =
    element) -> values ``[E, ...]``; ``idim`` lists reference dimensions to differentiate along (each once)."""
    d = function_space.dimension
    if len(x) != d:
        raise TypeError(f"reconstruct_batch() takes {d} coordinate arrays, {len(x)} were given.")
    xs = [np.ascontiguousarray(v, dtype=np.float64) for v in x]
    for v in xs[1:]:
        if v.shape != xs[0].shape:
            raise ValueError("All input coordinate arrays must have the same shape.")
    dof_values = np.ascontiguousarray(dof_values, dtype=np.float64)
    E = dof_values.shape[0]
    if xs[0].shape[:1] != (E,) or dof_values.size != E * function_space.n_basis:
        raise ValueError("dof_values must be [E, n0, n1, ..] and every coordinate array [E, ...].")
    mask = 0
    for k in idim:
        if not 0 <= int(k) < d or (mask >> int(k)) & 1:
            raise ValueError(f"Invalid or repeated derivative dimension {k}.")
        mask |= 1 << int(k)
    pts = np.stack([v.reshape(E, -1) for v in xs], axis=1)  # [E][d][P]
    out = _ctx().reconstruct_points(function_space.basis, dof_values.reshape(E, -1), pts, [mask])
    return out[:, 0, :].reshape(xs[0].shape)


def compute_dual_degrees_of_freedom(fn: Callable, integration_space: IntegrationSpace, function_space: FunctionSpace, /) -> DegreesOfFreedom:
    """python/interplib/degrees_of_freedom.py:40-71 (same result as projection_l2_dual on the reference domain)."""
    return projection_l2_dual(fn, function_space, integration_space)


# ------------------------------------------------------------------------------------------------ src/polynomials
def _lagrange_samples(roots, x, out, deriv: int) -> np.ndarray:
    """Shared body of lagrange1d / dlagrange1d (src/python/module.c:102-199, 307-404): ``roots`` 1-D, ``x`` any shape,
    result ``x.shape + (len(roots),)``; an ``out`` array is checked like the extension does and returned itself."""
    roots = np.ascontiguousarray(roots, dtype=np.float64)
    if roots.ndim != 1:
        raise ValueError("Roots must be a one-dimensional array.")
    x = np.ascontiguousarray(x, dtype=np.float64)
    shape = x.shape + (roots.size,)
    if out is not None:
        if not isinstance(out, np.ndarray):
            raise TypeError(f"out must be a numpy array, not {type(out).__name__}")
        if out.dtype != np.float64:
            raise ValueError("Output array must have the correct type (numpy.double/numpy.float64).")
        if out.shape != shape:
            raise ValueError("Output must have same shape as input array, except for one more dimension, which must be "
                             "the same length as the array of Lagrange nodes.")
    res = batch.lagrange_samples(roots, x.reshape(-1), deriv).reshape(shape)
    if out is None:
        return res
    out[...] = res
    return out


def lagrange1d(roots, x, out=None, /) -> np.ndarray:
    """Lagrange polynomials with nodes ``roots`` at ``x`` (src/python/module.c:102-199)."""
    return _lagrange_samples(roots, x, out, 0)


def dlagrange1d(roots, x, out=None, /) -> np.ndarray:
    """First derivatives of the Lagrange polynomials (src/python/module.c:307-404)."""
    return _lagrange_samples(roots, x, out, 1)


def d2lagrange1d(x, xp, /) -> np.ndarray:
    """Second derivatives, ``[len(x), len(xp)]``; note the argument order: points first, nodes second, both 1-D float64
    arrays (src/python/module.c:514-557)."""
    if not isinstance(x, np.ndarray) or not isinstance(xp, np.ndarray):
        raise TypeError("d2lagrange1d() arguments must be numpy arrays")
    if x.dtype != np.float64 or xp.dtype != np.float64 or x.ndim != 1 or xp.ndim != 1:
        raise ValueError("x and xp must be one-dimensional float64 arrays")
    return batch.lagrange_samples(np.ascontiguousarray(xp), np.ascontiguousarray(x), 2)


def bernstein1d(n: int, x, /) -> np.ndarray:
    """Bernstein polynomials of degree ``n`` on the unit interval, ``[len(x), n+1]`` (src/python/module.c:646-686)."""
    if int(n) < 0:
        raise OverflowError("can't convert negative value to unsigned int")
    x = np.ascontiguousarray(x, dtype=np.float64)
    if x.ndim != 1:
        raise ValueError("x must be a one-dimensional array")
    return batch.bernstein_samples(int(n), x)


def bernstein_coefficients(x, /) -> np.ndarray:
    """Power-series coefficients -> Bernstein coefficients of the same polynomial (src/python/module.c:706-718)."""
    c = np.array(x, dtype=np.float64, order="C")  # always a copy, like NPY_ARRAY_ENSURECOPY
    if c.ndim != 1:
        raise ValueError("x must be a one-dimensional array")
    return batch.bernstein_from_power_series(c)


def compute_gll(order: int, max_iter: int = 10, tol: float = 1e-15) -> tuple[np.ndarray, np.ndarray]:
    """Gauss-Legendre-Lobatto nodes and weights (src/python/module.c:559-621).  A node that does not converge raises
    RuntimeWarning, as the extension does (it sets the exception rather than issuing a warning)."""
    order, max_iter, tol = int(order), int(max_iter), float(tol)
    if order < 0:
        raise ValueError(f"Order must be positive, but was given as {order}.")
    if max_iter < 0:
        raise ValueError(f"Number of maximum iterations must be positive, but was given as {max_iter}.")
    if tol < 0:
        raise ValueError(f"Tolerance must be positive {tol:g}")
    nodes, weights, bad = batch.gauss_lobatto(order, max_iter, tol)
    if bad:
        raise RuntimeWarning(f"A total of {bad} nodes were non-converged. Consider changing the tolerance or increase the "
                             "number of iterations.")
    return nodes, weights


__all__ += ["lagrange1d", "dlagrange1d", "d2lagrange1d", "bernstein1d", "bernstein_coefficients", "compute_gll"]


def _sample_matrix(x, xp, deriv: int) -> np.ndarray:
    """``[N, M]`` matrix taking samples ``f(xp)`` to the ``deriv``-th derivative of their interpolant at the flattened
    ``x`` (python/interplib/lagrange.py:18-101)."""
    flat = np.ravel(np.atleast_1d(np.asarray(x, np.float64)), order="C")
    nodes = np.asarray(xp, np.float64)
    return d2lagrange1d(np.ascontiguousarray(flat), np.ascontiguousarray(nodes)) if deriv == 2 else _lagrange_samples(nodes, flat, None, deriv)


def lagrange_function_samples(x, xp) -> np.ndarray:
    return _sample_matrix(x, xp, 0)


def lagrange_derivative_samples(x, xp) -> np.ndarray:
    return _sample_matrix(x, xp, 1)


def lagrange_2derivative_samples(x, xp) -> np.ndarray:
    return _sample_matrix(x, xp, 2)


def _interp1d(x, xp, yp, deriv: int) -> np.ndarray:
    """python/interplib/lagrange.py:104-205: interpolant (or its derivative) through ``(xp, yp)`` evaluated at ``x``."""
    x, xp, yp = np.asarray(x, np.float64), np.asarray(xp, np.float64), np.asarray(yp, np.float64)
    if xp.ndim != 1 or yp.ndim != 1 or xp.shape != yp.shape:
        raise ValueError("Both xp and yp must be flat arrays of equal length")
    return np.reshape(_sample_matrix(x, xp, deriv) @ yp, x.shape).astype(np.float64, copy=False)


def interp1d_function_samples(x, xp, yp) -> np.ndarray:
    return _interp1d(x, xp, yp, 0)


def interp1d_derivative_samples(x, xp, yp) -> np.ndarray:
    return _interp1d(x, xp, yp, 1)


def interp1d_2derivative_samples(x, xp, yp) -> np.ndarray:
    return _interp1d(x, xp, yp, 2)


__all__ += ["lagrange_function_samples", "lagrange_derivative_samples", "lagrange_2derivative_samples",
            "interp1d_function_samples", "interp1d_derivative_samples", "interp1d_2derivative_samples"]


# ------------------------------------------------------------------------------------------------ geometry authoring
class HypercubeDomain:
    """An element described by one DoF set per physical coordinate; calling it with an IntegrationSpace builds
    the SpaceMap (python/interplib/domains.py:28-199)."""

    def __init__(self, *dofs: DegreesOfFreedom):
        if not dofs:
            raise ValueError("At least one coordinate must have its DoFs specified.")
        for i, d in enumerate(dofs):
            if not isinstance(d, DegreesOfFreedom):
                raise TypeError(f"Argument {i} was not {DegreesOfFreedom}, but {type(d)}.")
            if d.function_space != dofs[0].function_space:
                raise ValueError(f"Function spaces of the DoFs {i} does not match the rest!")
        self.dofs = tuple(dofs)

    ndim_physical = property(lambda self: len(self.dofs))
    ndim_reference = property(lambda self: len(self.dofs[0].shape))
    function_space = property(lambda self: self.dofs[0].function_space)

    def __call__(self, space: IntegrationSpace, /, *, integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                 basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY) -> SpaceMap:
        return SpaceMap(*(CoordinateMap(d, space, integration_registry, basis_registry) for d in self.dofs))

    def __eq__(self, other):
        return isinstance(other, HypercubeDomain) and len(self.dofs) == len(other.dofs) and all(
            a.function_space == b.function_space and np.array_equal(a.values, b.values) for a, b in zip(self.dofs, other.dofs))

    __hash__ = None

    @property
    def geometry_dofs(self) -> np.ndarray:
        """[n_phys][N_g]: one row of the batched API's ``geom_dofs[E][n_phys][N_g]``."""
        return np.stack([d.values.reshape(-1) for d in self.dofs])

    @property
    def endpoints(self) -> tuple[np.ndarray, ...]:
        """Coordinates of the 2^d corners: every coordinate reconstructed on the 2-point Lobatto rule per dimension."""
        corners = IntegrationSpace(*(IntegrationSpecs(1, "gauss-lobatto") for _ in range(self.ndim_reference)))
        return tuple(d.reconstruct_at_integration_points(corners) for d in self.dofs)

    def compute_size(self, int_space: IntegrationSpace | None = None, *,
                     integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                     basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY) -> float:
        """Length / area / volume: sum of weight * det J over a rule that integrates the Jacobian exactly."""
        if int_space is None:
            int_space = IntegrationSpace(*(IntegrationSpecs((order + 1) // 2) for order in self.function_space.orders))
        smap = self(int_space, integration_registry=integration_registry, basis_registry=basis_registry)
        return float(np.sum(int_space.weights(integration_registry) * smap.determinant))

    def sample(self, *x) -> tuple[np.ndarray, ...]:
        return tuple(reconstruct(d, *x) for d in self.dofs)

    def integrate(self, fn: Callable, int_space: IntegrationSpace, *,
                  integration_registry: IntegrationRegistry = DEFAULT_INTEGRATION_REGISTRY,
                  basis_registry: BasisRegistry = DEFAULT_BASIS_REGISTRY) -> float:
        return integrate_callable(fn, self(int_space, integration_registry=integration_registry, basis_registry=basis_registry),
                                  registry=integration_registry)

    def subregion(self, *ranges: tuple[float, float]) -> "HypercubeDomain":
        """The part of the domain whose reference coordinates lie in the given (low, high) pairs (missing trailing
        pairs = the whole [-1, 1]), re-interpolated on uniform Lagrange nodes of the same DoF shape."""
        d = self.ndim_reference
        if len(ranges) < d:  # reference quirk kept: fewer pairs than dimensions is the error, despite the wording
            raise ValueError(f"At most {d} pairs of divisions can be specified.")
        limits = [(float(lo), float(hi)) for lo, hi in ranges] + [(-1.0, +1.0)] * max(0, d - len(ranges))
        shape = self.dofs[0].shape
        if len(limits) != len(shape):
            raise ValueError("zip() argument lengths differ")
        grid = np.meshgrid(*(np.linspace(lo, hi, n) for n, (lo, hi) in zip(shape, limits)), indexing="ij")
        fs = FunctionSpace(*(BasisSpecs(BasisType.LAGRANGE_UNIFORM, n - 1) for n in shape))
        return HypercubeDomain(*(DegreesOfFreedom(fs, vals) for vals in self.sample(*grid)))


class Line(HypercubeDomain):
    """Curve through control points: Bernstein DoFs of order ``len(knots) - 1`` per coordinate
    (python/interplib/domains.py:200-233)."""

    def __init__(self, *knots):
        pts = np.array(knots)
        if pts.ndim != 2:
            raise ValueError("Line must be specified by an array of points.")
        if pts.shape[0] < 2:
            raise ValueError("At least two points must be given for a line.")
        fs = FunctionSpace(BasisSpecs(BasisType.BERNSTEIN, pts.shape[0] - 1))
        self.knots = pts
        super().__init__(*(DegreesOfFreedom(fs, pts[:, k]) for k in range(pts.shape[1])))

    start = property(lambda self: self.knots[0, :])
    end = property(lambda self: self.knots[-1, :])

    def reverse(self) -> "Line":
        return Line(*self.knots[::-1])


class Quad(HypercubeDomain):
    """Quadrilateral bounded by four :class:`Line` s given counter-clockwise (bottom, right, top, left), each starting
    where the previous one ends (python/interplib/domains.py:236-331).  Its coordinates are the transfinite (Coons)
    interpolant of the edges on uniform Lagrange nodes: edge contributions blended linearly across the opposite
    direction, minus the bilinear interpolant of the corners, which the two blends count twice."""

    def __init__(self, bottom: Line, right: Line, top: Line, left: Line):
        edges = (bottom, right, top, left)
        for ln in edges:
            if type(ln) is not Line:
                raise TypeError(f"Only {Line} objects can be used as inputs for a Quad")
        for prev, nxt, msg in ((bottom, right, "The right side does not start where the bottom ends."),
                               (right, top, "The top side does not start where the right ends."),
                               (top, left, "The left side does not start where the top ends."),
                               (left, bottom, "The bottom side does not start where the left ends.")):
            if np.any(prev.end != nxt.start):
                raise ValueError(msg)
        n_h = max(1, bottom.function_space.orders[0], top.function_space.orders[0])
        n_v = max(1, right.function_space.orders[0], left.function_space.orders[0])
        fs = FunctionSpace(BasisSpecs(BasisType.LAGRANGE_UNIFORM, n_h), BasisSpecs(BasisType.LAGRANGE_UNIFORM, n_v))
        xi, eta = np.linspace(-1, +1, n_h + 1), np.linspace(-1, +1, n_v + 1)
        # the top and left edges run against their reference direction
        on_b, on_r, on_t, on_l = bottom.sample(xi), right.sample(eta), top.sample(xi[::-1]), left.sample(eta[::-1])
        # blending weights on the [eta][xi] grid, written like the reference does ((1 -/+ s) / 2) so that the DoFs
        # round the same way
        lo_x, hi_x = ((1 - xi) / 2)[None, :], ((1 + xi) / 2)[None, :]
        lo_y, hi_y = ((1 - eta) / 2)[:, None], ((1 + eta) / 2)[:, None]
        dofs = []
        for b, r, t, l, c_bl, c_br, c_tr, c_tl in zip(on_b, on_r, on_t, on_l, bottom.start, right.start, top.start, left.start):
            edges_part = b[None, :] * lo_y + r[:, None] * hi_x + t[None, :] * hi_y + l[:, None] * lo_x
            corner_part = c_bl * lo_y * lo_x + c_br * lo_y * hi_x + c_tr * hi_y * hi_x + c_tl * hi_y * lo_x
            dofs.append(DegreesOfFreedom(fs, (edges_part - corner_part).T))
        super().__init__(*dofs)

    @classmethod
    def from_corners(cls, bottom_left, bottom_right, top_right, top_left) -> "Quad":
        """Straight-sided quad through four corners."""
        return cls(bottom=Line(bottom_left, bottom_right), right=Line(bottom_right, top_right),
                   top=Line(top_right, top_left), left=Line(top_left, bottom_left))


__all__ += ["HypercubeDomain", "Line", "Quad"]
