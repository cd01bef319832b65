"""/root/reference/python/interplib/integration.py"""
from interplib_b200.interp import integrate_callable, projection_l2_dual, projection_l2_primal  # noqa: F401
