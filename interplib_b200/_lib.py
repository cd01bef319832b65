"""ctypes binding of ``libinterp_b200.so`` (the C ABI declared in ``include/interp_b200.h``).

The shared library is the product: hand-written CUDA for sm_100a plus the C host-table layer.  This
module only loads it and declares the prototypes; it never substitutes a Python/NumPy implementation.
If the library is missing the import fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("INTERP_B200_LIB", _HERE / "libinterp_b200.so"))


class InterpError(RuntimeError):
    """A C-ABI call returned a non-zero ``interp_result_t``."""

    def __init__(self, code: int, name: str, message: str, detail: str):
        self.code = code
        self.name = name
        super().__init__(f"{name} ({message})" + (f": {detail}" if detail else ""))


class BasisSpec(C.Structure):
    """``basis_spec_t`` (reference src/basis/basis_set.h:22-26)."""

    _fields_ = [("type", C.c_int), ("order", C.c_uint)]


class Covector(C.Structure):
    """``ib200_covector_t`` = ``covector_basis_t`` (reference src/basis/covector_basis.h:21-26): the C struct packs
    ``dimension : 31`` and ``sign : 1`` into its first word (GCC allocates bit fields from the least significant bit)."""

    _fields_ = [("dim_sign", C.c_uint), ("basis_bits", C.c_uint)]

    @classmethod
    def make(cls, dimension: int, sign: int, bits: int) -> "Covector":
        return cls((dimension & 0x7FFFFFFF) | ((1 if sign else 0) << 31), bits)

    dimension = property(lambda self: self.dim_sign & 0x7FFFFFFF)
    sign = property(lambda self: self.dim_sign >> 31)


class IntegrationSpec(C.Structure):
    """``integration_spec_t`` (reference src/integration/integration_rules.h:17-21)."""

    _fields_ = [("type", C.c_int), ("order", C.c_uint)]


# enum numbering of the reference (src/basis/basis_set.h:11-20, src/integration/integration_rules.h:7-12)
BASIS_TYPES = {
    "legendre": 1,
    "lagrange-gauss-lobatto": 2,
    "lagrange-gauss": 3,
    "lagrange-uniform": 4,
    "lagrange-chebyshev-gauss": 5,
    "bernstein": 6,
}
BASIS_NAMES = {v: k for k, v in BASIS_TYPES.items()}
RULE_TYPES = {"gauss": 1, "gauss-lobatto": 2}
RULE_NAMES = {v: k for k, v in RULE_TYPES.items()}

HOST, DEVICE = 0, 1

# error codes that map to ValueError / MemoryError like the reference extension does
_VALUE_ERRORS = {5, 102}
COMM_ID_BYTES = 128
_MEMORY_ERRORS = {3}

_dp = C.POINTER(C.c_double)
_PROTOTYPES = {
    # name: (restype, argtypes)
    "ib200_error_str": (C.c_char_p, [C.c_int]),
    "ib200_error_msg": (C.c_char_p, [C.c_int]),
    "ib200_last_error": (C.c_char_p, []),
    "ib200_version": (C.c_char_p, []),
    "ib200_rule_1d": (C.c_int, [IntegrationSpec, _dp, _dp, C.POINTER(C.c_uint)]),
    "ib200_rule_order_for_accuracy": (C.c_int, [C.c_int, C.c_uint, C.POINTER(C.c_uint)]),
    "ib200_lagrange_roots": (C.c_int, [BasisSpec, _dp]),
    "ib200_basis_table_1d": (C.c_int, [BasisSpec, IntegrationSpec, _dp, _dp]),
    "ib200_integration_space_weights": (C.c_int, [C.c_uint, C.POINTER(IntegrationSpec), _dp]),
    "ib200_integration_space_nodes": (C.c_int, [C.c_uint, C.POINTER(IntegrationSpec), _dp]),
    "ib200_values_at_integration_nodes": (C.c_int, [C.c_uint, C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), _dp]),
    "ib200_basis_at_points": (C.c_int, [BasisSpec, C.c_size_t, _dp, _dp, _dp]),
    "ib200_function_space_evaluate": (C.c_int, [C.c_uint, C.POINTER(BasisSpec), C.c_size_t, C.POINTER(_dp), _dp]),
    "ib200_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "ib200_ctx_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "ib200_ctx_destroy": (C.c_int, [C.c_void_p]),
    "ib200_ctx_synchronize": (C.c_int, [C.c_void_p]),
    "ib200_ctx_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "ib200_ctx_info": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_size_t)]),
    "ib200_ctx_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_ulonglong), C.POINTER(C.c_uint)]),
    "ib200_ctx_clear_cache": (C.c_int, [C.c_void_p]),
    "ib200_ctx_set_force_generic": (C.c_int, [C.c_void_p, C.c_int]),
    "ib200_ctx_set_mma_min_points": (C.c_int, [C.c_void_p, C.c_int]),
    "ib200_ctx_set_lo_kernel": (C.c_int, [C.c_void_p, C.c_int]),
    "ib200_ctx_set_fastdiag": (C.c_int, [C.c_void_p, C.c_int]),
    "ib200_ctx_set_kron": (C.c_int, [C.c_void_p, C.c_int]),
    "ib200_ctx_set_kron_gradient": (C.c_int, [C.c_void_p, C.c_int]),
    "ib200_ctx_set_chunk_elements": (C.c_int, [C.c_void_p, C.c_size_t]),
    "ib200_device_alloc": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "ib200_device_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "ib200_pinned_alloc": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "ib200_pinned_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "ib200_copy_h2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "ib200_copy_d2h": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "ib200_device_memset": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_size_t]),
    "ib200_timer_start": (C.c_int, [C.c_void_p, C.c_uint]),
    "ib200_timer_stop": (C.c_int, [C.c_void_p, C.c_uint]),
    "ib200_timer_elapsed_ms": (C.c_int, [C.c_void_p, C.c_uint, C.POINTER(C.c_float)]),
    "ib200_reconstruct_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), C.c_size_t, C.c_void_p, C.c_uint,
         C.POINTER(C.c_uint), C.c_void_p, C.c_int],
    ),
    "ib200_evaluate_points_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int],
    ),
    "ib200_reconstruct_points_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint,
         C.POINTER(C.c_uint), C.c_void_p, C.c_int],
    ),
    "ib200_incidence_matrix": (C.c_int, [BasisSpec, C.c_void_p]),
    "ib200_pair_moments": (C.c_int, [BasisSpec, BasisSpec, IntegrationSpec, C.c_uint, C.c_int, C.c_void_p]),
    "ib200_fastdiag_tables": (C.c_int, [BasisSpec, IntegrationSpec, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ib200_dof_derivative_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.c_uint, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int],
    ),
    "ib200_covector_make": (C.c_int, [C.c_uint, C.c_int, C.c_uint, C.POINTER(C.c_uint), C.POINTER(Covector)]),
    "ib200_covector_rank": (C.c_uint, [Covector]),
    "ib200_covector_is_zero": (C.c_int, [Covector]),
    "ib200_covector_wedge": (C.c_int, [Covector, Covector, C.POINTER(Covector)]),
    "ib200_covector_hodge": (C.c_int, [Covector, C.POINTER(Covector)]),
    "ib200_covector_apply_contra": (C.c_int, [Covector, C.c_uint, C.POINTER(Covector)]),
    "ib200_covector_order": (C.c_int, [Covector, Covector]),
    "ib200_kform_component_count": (C.c_uint, [C.c_uint, C.c_uint]),
    "ib200_kform_components": (C.c_int, [C.c_uint, C.c_uint, C.POINTER(Covector)]),
    "ib200_kform_component_space": (C.c_int, [C.c_uint, C.POINTER(BasisSpec), Covector, C.POINTER(BasisSpec)]),
    "ib200_kform_dof_count": (C.c_int, [C.c_uint, C.POINTER(BasisSpec), C.c_uint, C.POINTER(C.c_size_t)]),
    "ib200_dof_offsets": (C.c_int, [C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "ib200_geometry_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.c_uint, C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), C.c_size_t, C.c_void_p,
         C.c_void_p, C.c_void_p, C.c_void_p, C.c_int],
    ),
    "ib200_mass_matrix_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), C.c_size_t,
         C.c_void_p, C.c_int, C.c_void_p, C.c_int],
    ),
    "ib200_mass_matrix_mapped_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), C.c_uint,
         C.POINTER(BasisSpec), C.c_size_t, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int],
    ),
    "ib200_l2_dual_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), C.c_size_t, C.c_void_p, C.c_void_p,
         C.c_void_p, C.c_int],
    ),
    "ib200_l2_dual_mapped_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), C.c_uint, C.POINTER(BasisSpec),
         C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int],
    ),
    "ib200_l2_system_mapped_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), C.c_uint, C.POINTER(BasisSpec),
         C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int],
    ),
    "ib200_kform_mass_mapped_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.c_uint, C.POINTER(IntegrationSpec), C.c_uint, C.POINTER(BasisSpec),
         C.c_size_t, C.c_void_p, C.c_void_p, C.c_int],
    ),
    "ib200_l2_project_mapped_batch": (
        C.c_int,
        [C.c_void_p, C.c_uint, C.POINTER(BasisSpec), C.POINTER(IntegrationSpec), C.c_uint, C.POINTER(BasisSpec),
         C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int],
    ),
    "ib200_spd_solve_batch": (C.c_int, [C.c_void_p, C.c_uint, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]),
    "ib200_dot": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "ib200_dot_static": (C.c_int, [C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "ib200_lagrange_samples": (C.c_int, [C.c_uint, C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint, C.c_void_p]),
    "ib200_gauss_lobatto": (C.c_int, [C.c_uint, C.c_uint, C.c_double, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]),
    "ib200_bernstein_samples": (C.c_int, [C.c_uint, C.c_size_t, C.c_void_p, C.c_void_p]),
    "ib200_bernstein_from_power_series": (C.c_int, [C.c_uint, C.c_void_p]),
    "ib200_comm_unique_id": (C.c_int, [C.c_void_p]),
    "ib200_comm_create": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "ib200_comm_destroy": (C.c_int, [C.c_void_p]),
    "ib200_comm_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "ib200_comm_check": (C.c_int, [C.c_void_p]),
    "ib200_partition_range": (C.c_int, [C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "ib200_allgather_slabs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p]),
    "ib200_allgatherv": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), C.c_int]),
    "ib200_comm_join": (C.c_int, [C.c_void_p]),
    "ib200_comm_register_gather": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "ib200_comm_unregister_gather": (C.c_int, [C.c_void_p]),
    "ib200_allreduce_sums": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
}

EXPORTED_SYMBOLS = tuple(_PROTOTYPES)


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C interplib_b200/csrc`).  There is no pure-Python fallback."
        )
    lib = C.CDLL(str(LIB_PATH))
    for name, (restype, argtypes) in _PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    return lib


lib = _load()


def check(code: int) -> None:
    """Raise on a non-zero result, mirroring the reference extension's exception mapping
    (RuntimeError / ValueError / MemoryError, e.g. src/python/integration_objects.c:910)."""
    if code == 0:
        return
    name = lib.ib200_error_str(code).decode()
    msg = lib.ib200_error_msg(code).decode()
    detail = lib.ib200_last_error().decode()
    err = InterpError(code, name, msg, detail)
    if code in _VALUE_ERRORS:
        raise ValueError(str(err)) from err
    if code in _MEMORY_ERRORS:
        raise MemoryError(str(err)) from err
    raise err


def basis_specs(specs) -> C.Array:
    """[(type_name | int, order), ...] -> basis_spec_t[]"""
    arr = (BasisSpec * len(specs))()
    for i, (t, o) in enumerate(specs):
        arr[i].type = BASIS_TYPES[t] if isinstance(t, str) else int(t)
        arr[i].order = int(o)
    return arr


def rule_specs(specs) -> C.Array:
    """[(method_name | int, order), ...] -> integration_spec_t[]"""
    arr = (IntegrationSpec * len(specs))()
    for i, (t, o) in enumerate(specs):
        arr[i].type = RULE_TYPES[t] if isinstance(t, str) else int(t)
        arr[i].order = int(o)
    return arr
