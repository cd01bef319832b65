"""Stand-in for the reference's CPython extension ``interplib._interp`` (module table and type list:
/root/reference/src/python/module.c:730-898): the same names, backed by ``libinterp_b200.so`` through
``interplib_b200.interp`` / ``interplib_b200.kform``.  The reference's own ``python/interplib/*.py`` run unchanged on top
of this module (tests/test_dropin_gpu.py does exactly that)."""
import os
import sys

_repo = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _repo not in sys.path:  # the alias directory alone on the path is enough
    sys.path.append(_repo)

from interplib_b200.interp import (  # noqa: E402,F401
    DEFAULT_BASIS_REGISTRY,
    DEFAULT_INTEGRATION_REGISTRY,
    BasisRegistry,
    BasisSpecs,
    CoordinateMap,
    DegreesOfFreedom,
    FunctionSpace,
    IntegrationRegistry,
    IntegrationSpace,
    IntegrationSpecs,
    SpaceMap,
    bernstein1d,
    bernstein_coefficients,
    compute_gll,
    compute_gradient_mass_matrix,
    compute_mass_matrix,
    d2lagrange1d,
    dlagrange1d,
    incidence_matrix,
    lagrange1d,
)
from interplib_b200.kform import CovectorBasis  # noqa: E402,F401
