"""/root/reference/python/interplib/lagrange.py"""
from interplib_b200.interp import (  # noqa: F401
    interp1d_2derivative_samples,
    interp1d_derivative_samples,
    interp1d_function_samples,
    lagrange_2derivative_samples,
    lagrange_derivative_samples,
    lagrange_function_samples,
)
