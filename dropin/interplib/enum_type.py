"""/root/reference/python/interplib/enum_type.py"""
from interplib_b200.interp import BasisType, IntegrationMethod  # noqa: F401
