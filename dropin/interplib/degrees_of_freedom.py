"""/root/reference/python/interplib/degrees_of_freedom.py"""
from interplib_b200.interp import compute_dual_degrees_of_freedom, reconstruct  # noqa: F401
