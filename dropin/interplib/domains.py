"""/root/reference/python/interplib/domains.py"""
from interplib_b200.interp import HypercubeDomain, Line, Quad  # noqa: F401
