"""``import interplib`` resolved to the B200 implementation.

Put this directory (``<repo>/dropin``) on ``sys.path`` / ``PYTHONPATH`` ahead of an installed reference and existing
user code keeps working: the same names as /root/reference/python/interplib/__init__.py:3-50, every one re-exported
from ``interplib_b200`` (numeric calls run on the GPU through ``libinterp_b200.so``; there is no CPU fallback).
The topology types of the reference (GeoID, Line, Surface, Manifold*) are outside the hot path and are not provided.
"""
from interplib._interp import DEFAULT_BASIS_REGISTRY as DEFAULT_BASIS_REGISTRY
from interplib._interp import DEFAULT_INTEGRATION_REGISTRY as DEFAULT_INTEGRATION_REGISTRY
from interplib._interp import BasisRegistry as BasisRegistry
from interplib._interp import BasisSpecs as BasisSpecs
from interplib._interp import CoordinateMap as CoordinateMap
from interplib._interp import DegreesOfFreedom as DegreesOfFreedom
from interplib._interp import FunctionSpace as FunctionSpace
from interplib._interp import IntegrationRegistry as IntegrationRegistry
from interplib._interp import IntegrationSpace as IntegrationSpace
from interplib._interp import IntegrationSpecs as IntegrationSpecs
from interplib._interp import bernstein1d as bernstein1d
from interplib._interp import bernstein_coefficients as bernstein_coefficients
from interplib._interp import compute_gll as compute_gll
from interplib._interp import compute_mass_matrix as compute_mass_matrix
from interplib._interp import dlagrange1d as dlagrange1d
from interplib._interp import lagrange1d as lagrange1d
from interplib.degrees_of_freedom import compute_dual_degrees_of_freedom as compute_dual_degrees_of_freedom
from interplib.degrees_of_freedom import reconstruct as reconstruct
from interplib.domains import Line as Line
from interplib.domains import Quad as Quad
from interplib.enum_type import BasisType as BasisType
from interplib.enum_type import IntegrationMethod as IntegrationMethod
from interplib.integration import integrate_callable as integrate_callable
from interplib.integration import projection_l2_dual as projection_l2_dual
from interplib.integration import projection_l2_primal as projection_l2_primal
from interplib.lagrange import interp1d_2derivative_samples as interp1d_2derivative_samples
from interplib.lagrange import interp1d_derivative_samples as interp1d_derivative_samples
from interplib.lagrange import interp1d_function_samples as interp1d_function_samples
from interplib.lagrange import lagrange_2derivative_samples as lagrange_2derivative_samples
from interplib.lagrange import lagrange_derivative_samples as lagrange_derivative_samples
from interplib.lagrange import lagrange_function_samples as lagrange_function_samples
