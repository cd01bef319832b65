"""Element-batched front-end of the C ABI: NumPy arrays in, NumPy arrays out, the GPU in between.

The reference has no batched API (one element per Python call); these functions
are the batched counterparts of its per-element entry points and call straight into
``libinterp_b200.so`` with host pointers.  ``Context.device_*`` variants keep data resident on the GPU.

Specs are plain tuples: a function space is ``[(basis_type, order), ...]`` (one entry per reference
dimension, names as in ``interplib.BasisType``) and an integration space is ``[(method, order), ...]``.
"""
from __future__ import annotations

import ctypes as C
import threading
import weakref

import numpy as np

from . import _lib
from ._lib import DEVICE, HOST, basis_specs, check, lib, rule_specs


def _n_basis(basis) -> int:
    return int(np.prod([int(o) + 1 for _, o in basis], dtype=np.int64))


def _n_points(rules) -> int:
    return int(np.prod([int(o) + 1 for _, o in rules], dtype=np.int64))


def _as_in(a, shape) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a.reshape(shape)


def _as_out(out, shape, name="out") -> np.ndarray:
    """A fresh output array, or the caller's after the checks the reference applies to its ``out=`` arguments
    (``check_input_array`` with NPY_DOUBLE / C_CONTIGUOUS / ALIGNED, e.g. src/python/degrees_of_freedom.c:377): the C
    side writes ``prod(shape)`` doubles through the raw pointer, so anything else would corrupt the caller's heap."""
    shape = tuple(int(v) for v in shape)
    if out is None:
        return np.empty(shape)
    if not isinstance(out, np.ndarray):
        raise TypeError(f"{name} must be a NumPy array (or a device pointer with where=DEVICE), got {type(out).__name__}")
    if out.dtype != np.float64 or not out.flags.c_contiguous or not out.flags.writeable or not out.flags.aligned:
        raise ValueError(f"{name} must be a writeable, aligned, C-contiguous float64 array")
    if out.size != int(np.prod(shape, dtype=np.int64)):
        raise ValueError(f"{name} must hold {int(np.prod(shape, dtype=np.int64))} values (shape {shape}), it holds {out.size}")
    return out


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))  # raw device / pinned address, DeviceBuffer, DeviceArray


def _release_device_memory(ctx: "Context", ptr: int) -> None:
    """Finaliser of an owned :class:`DeviceBuffer`: runs when the last reference is dropped.  A context that is already closed
    released the memory itself (``Context.close``)."""
    if ctx.handle and ptr:
        lib.ib200_device_free(ctx.handle, ptr)


class DeviceBuffer:
    """A device allocation owned by a :class:`Context` (``ib200_device_alloc``).

    ``owned=False`` (``Context.alloc``): raw memory whose address the caller passes around as an integer; it lives until
    :meth:`free` or the end of the process.  ``owned=True`` (every buffer behind a :class:`DeviceArray`): the memory is
    returned when the last Python reference to the buffer goes away -- the arrays that share it (``reshape`` views) each
    hold one -- or when the context is closed, whichever comes first."""

    def __init__(self, ctx: "Context", nbytes: int, owned: bool = False):
        self.ctx = ctx
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        check(lib.ib200_device_alloc(ctx.handle, self.nbytes, C.byref(p)))
        self.ptr = p.value
        self._finalizer = None
        if owned:
            self._finalizer = weakref.finalize(self, _release_device_memory, ctx, self.ptr)
            self._finalizer.atexit = False
            ctx._owned.add(self)

    def upload(self, host: np.ndarray) -> "DeviceBuffer":
        host = np.ascontiguousarray(host)
        assert host.nbytes <= self.nbytes
        check(lib.ib200_copy_h2d(self.ctx.handle, self.ptr, host.ctypes.data_as(C.c_void_p), host.nbytes))
        return self

    def download(self, shape, dtype=np.float64, offset_bytes: int = 0) -> np.ndarray:
        out = np.empty(shape, dtype=dtype)
        assert offset_bytes + out.nbytes <= self.nbytes
        check(lib.ib200_copy_d2h(self.ctx.handle, out.ctypes.data_as(C.c_void_p), self.ptr + offset_bytes, out.nbytes))
        return out

    def free(self) -> None:
        if self.ptr:
            if self._finalizer is not None:
                self._finalizer.detach()
            check(lib.ib200_device_free(self.ctx.handle, self.ptr))
            self.ptr = None

    def __int__(self) -> int:
        if not self.ptr:
            raise ValueError("device buffer has been freed")
        return int(self.ptr)


class DeviceArray:
    """A float64 array that lives in device memory: what the batched calls return when their inputs are device arrays, so
    that chains (geometry + f -> primal DoFs -> values and gradients at the rule nodes -> ...) never cross PCIe.  Explicit
    transfers only: :meth:`Context.to_device`, :meth:`to_host`.  ``int(a)`` is the device address (usable wherever the
    ``where=DEVICE`` form of a call takes a raw pointer).

    Ownership: the memory belongs to the array and to every ``reshape`` view of it together and is released when the last of
    them is dropped (intermediates of a chain free themselves) or when the context closes.  :meth:`free` releases it at
    once for ALL views; using any of them afterwards raises."""

    def __init__(self, ctx: "Context", shape, buf: DeviceBuffer | None = None):
        self.ctx = ctx
        self.shape = tuple(int(v) for v in shape)
        self.size = int(np.prod(self.shape, dtype=np.int64))
        self.buf = buf if buf is not None else DeviceBuffer(ctx, max(self.size, 1) * 8, owned=True)
        assert self.buf.nbytes >= self.size * 8

    @property
    def ptr(self) -> int:
        return int(self.buf)

    def __int__(self) -> int:
        return int(self.buf)

    def reshape(self, *shape) -> "DeviceArray":
        """A view of the same memory with another shape (one ``-1`` allowed)."""
        shape = tuple(shape[0]) if len(shape) == 1 and isinstance(shape[0], (tuple, list)) else tuple(shape)
        known = int(np.prod([v for v in shape if v != -1], dtype=np.int64))
        shape = tuple(self.size // max(known, 1) if v == -1 else int(v) for v in shape)
        if int(np.prod(shape, dtype=np.int64)) != self.size:
            raise ValueError(f"cannot reshape a device array of {self.size} values to {shape}")
        return DeviceArray(self.ctx, shape, self.buf)

    def to_host(self) -> np.ndarray:
        self.ctx.synchronize()
        return self.buf.download(self.shape)

    def free(self) -> None:
        self.buf.free()


def _is_dev(*arrays) -> bool:
    return any(isinstance(a, DeviceArray) for a in arrays)


def _dev_in(a, shape, name="input") -> "DeviceArray":
    if not isinstance(a, DeviceArray):
        raise TypeError(f"{name}: device-resident calls take DeviceArray inputs throughout (Context.to_device), got {type(a).__name__}")
    return a.reshape(shape)


class PinnedArray:
    """Page-locked host memory viewed as a NumPy array (``ib200_pinned_alloc``)."""

    def __init__(self, ctx: "Context", shape, dtype=np.float64):
        self.ctx = ctx
        n = int(np.prod(shape, dtype=np.int64)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        check(lib.ib200_pinned_alloc(ctx.handle, n, C.byref(p)))
        self.ptr = p.value
        buf = (C.c_char * max(n, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape, dtype=np.int64))).reshape(shape)

    def free(self) -> None:
        if self.ptr:
            self.array = None
            check(lib.ib200_pinned_free(self.ctx.handle, self.ptr))
            self.ptr = None


class Context:
    """One GPU context (``ib200_ctx_create``).  Creation fails loudly when no B200 is visible."""

    def __init__(self, device: int = 0):
        h = C.c_void_p()
        check(lib.ib200_ctx_create(int(device), C.byref(h)))
        self.handle = h
        self.device = int(device)
        self._owned = weakref.WeakSet()  # DeviceArray buffers still alive: released with the context

    # -- housekeeping ---------------------------------------------------------------------------
    def close(self) -> None:
        if self.handle:
            for buf in list(self._owned):
                buf.free()
            check(lib.ib200_ctx_destroy(self.handle))
            self.handle = None

    def synchronize(self) -> None:
        check(lib.ib200_ctx_synchronize(self.handle))

    def set_stream(self, cuda_stream: int | None) -> None:
        check(lib.ib200_ctx_set_stream(self.handle, C.c_void_p(cuda_stream or 0)))

    def set_force_generic(self, on: bool) -> None:
        check(lib.ib200_ctx_set_force_generic(self.handle, int(bool(on))))

    def set_mma_min_points(self, points: int) -> None:
        """Quadrature points per dimension from which the DMMA mass kernel is used (0 = never)."""
        check(lib.ib200_ctx_set_mma_min_points(self.handle, int(points)))

    def set_lo_kernel(self, on: bool) -> None:
        """Compile-time specialised low-order mass kernel on/off (testing aid)."""
        check(lib.ib200_ctx_set_lo_kernel(self.handle, int(bool(on))))

    def set_kron(self, on: bool) -> None:
        """Kronecker-sum mass kernel for bilinear quads on/off (off: the sum-factorised DFMA / DMMA kernels)."""
        check(lib.ib200_ctx_set_kron(self.handle, int(bool(on))))

    def set_kron_gradient(self, on: bool) -> None:
        """Kronecker-sum kernel also for gradient mass matrices of bilinear quads (default off: see ib200_ctx_set_kron_gradient)."""
        check(lib.ib200_ctx_set_kron_gradient(self.handle, int(bool(on))))

    def set_fastdiag(self, on: bool) -> None:
        """Fast-diagonalisation projection of bilinear quads on/off (off: every mass matrix is assembled and factorised)."""
        check(lib.ib200_ctx_set_fastdiag(self.handle, int(bool(on))))

    def set_chunk_elements(self, n: int) -> None:
        check(lib.ib200_ctx_set_chunk_elements(self.handle, int(n)))

    def clear_cache(self) -> None:
        check(lib.ib200_ctx_clear_cache(self.handle))

    def info(self) -> dict:
        name = C.create_string_buffer(256)
        sm, cc, mem = C.c_int(), C.c_int(), C.c_size_t()
        check(lib.ib200_ctx_info(self.handle, name, 256, C.byref(sm), C.byref(cc), C.byref(mem)))
        return {"name": name.value.decode(), "sm_count": sm.value, "cc": cc.value, "mem_bytes": mem.value}

    def stats(self) -> dict:
        n, t = C.c_ulonglong(), C.c_uint()
        check(lib.ib200_ctx_stats(self.handle, C.byref(n), C.byref(t)))
        return {"kernel_launches": n.value, "cached_tables": t.value}

    def alloc(self, nbytes: int) -> DeviceBuffer:
        return DeviceBuffer(self, nbytes)

    def pinned(self, shape, dtype=np.float64) -> PinnedArray:
        return PinnedArray(self, shape, dtype)

    def to_device(self, host) -> DeviceArray:
        """Copy a NumPy array to the device; the batched calls keep their results there when given such inputs."""
        host = np.ascontiguousarray(host, dtype=np.float64)
        d = DeviceArray(self, host.shape)
        if host.size:
            d.buf.upload(host)
        return d

    def _dev_out(self, out, shape) -> DeviceArray:
        if out is None:
            return DeviceArray(self, shape)
        if not isinstance(out, DeviceArray) or out.size != int(np.prod(shape, dtype=np.int64)):
            raise ValueError(f"out must be a DeviceArray of {int(np.prod(shape, dtype=np.int64))} values")
        return out.reshape(shape)

    def timer_start(self, slot: int = 0) -> None:
        check(lib.ib200_timer_start(self.handle, slot))

    def timer_stop(self, slot: int = 0) -> None:
        check(lib.ib200_timer_stop(self.handle, slot))

    def timer_ms(self, slot: int = 0) -> float:
        ms = C.c_float()
        check(lib.ib200_timer_elapsed_ms(self.handle, slot, C.byref(ms)))
        return float(ms.value)

    # -- K1 ----------------------------------------------------------------------------------------
    def reconstruct(self, basis, rules, dofs, deriv_masks=(0,), out=None, where=HOST, n_elements=None):
        """dofs[E][N] -> out[E][n_fields][Q]   (ib200_reconstruct_batch)"""
        N, Q = _n_basis(basis), _n_points(rules)
        masks = (C.c_uint * len(deriv_masks))(*[int(m) for m in deriv_masks])
        if _is_dev(dofs, out):
            dofs = _dev_in(dofs, (-1, N), "dofs")
            where, n_elements = DEVICE, dofs.shape[0]
            out = self._dev_out(out, (n_elements, len(deriv_masks), Q))
        if where == HOST:
            dofs = _as_in(dofs, (-1, N))
            E = dofs.shape[0]
            out = _as_out(out, (E, len(deriv_masks), Q))
        else:
            E = int(n_elements)
        check(lib.ib200_reconstruct_batch(self.handle, len(basis), basis_specs(basis), rule_specs(rules), E, _ptr(dofs),
                                          len(deriv_masks), masks, _ptr(out), where))
        return out

    # -- N2 ----------------------------------------------------------------------------------------
    def evaluate_points(self, basis, x, out=None, where=HOST, n_elements=None, n_points=None):
        """x[E][ndim][P] -> out[E][P][N]: FunctionSpace.evaluate per element   (ib200_evaluate_points_batch)"""
        N, ndim = _n_basis(basis), len(basis)
        if where == HOST:
            x = np.ascontiguousarray(x, dtype=np.float64)
            assert x.ndim == 3 and x.shape[1] == ndim, "x must be [E][ndim][P]"
            E, P = x.shape[0], x.shape[2]
            out = _as_out(out, (E, P, N))
        else:
            E, P = int(n_elements), int(n_points)
        check(lib.ib200_evaluate_points_batch(self.handle, ndim, basis_specs(basis), E, P, _ptr(x), _ptr(out), where))
        return out

    def reconstruct_points(self, basis, dofs, x, deriv_masks=(0,), out=None, where=HOST, n_elements=None, n_points=None):
        """dofs[E][N], x[E][ndim][P] -> out[E][n_fields][P]: reconstruct() per element   (ib200_reconstruct_points_batch)"""
        N, ndim = _n_basis(basis), len(basis)
        masks = (C.c_uint * len(deriv_masks))(*[int(m) for m in deriv_masks])
        if where == HOST:
            dofs = _as_in(dofs, (-1, N))
            E = dofs.shape[0]
            x = np.ascontiguousarray(x, dtype=np.float64)
            assert x.ndim == 3 and x.shape[0] == E and x.shape[1] == ndim, "x must be [E][ndim][P]"
            P = x.shape[2]
            out = _as_out(out, (E, len(deriv_masks), P))
        else:
            E, P = int(n_elements), int(n_points)
        check(lib.ib200_reconstruct_points_batch(self.handle, ndim, basis_specs(basis), E, _ptr(dofs), P, _ptr(x),
                                                 len(deriv_masks), masks, _ptr(out), where))
        return out

    # -- N3 ----------------------------------------------------------------------------------------
    def dof_derivative(self, basis, dofs, idim, out=None, where=HOST, n_elements=None):
        """dofs[E][N] -> DoFs of d/dxi_idim in the space one order lower along idim   (ib200_dof_derivative_batch)"""
        N = _n_basis(basis)
        lower = [(t, o - 1 if k == idim else o) for k, (t, o) in enumerate(basis)]
        if _is_dev(dofs, out):
            dofs = _dev_in(dofs, (-1, N), "dofs")
            where, n_elements = DEVICE, dofs.shape[0]
            out = self._dev_out(out, (n_elements, _n_basis(lower) if 0 <= idim < len(basis) and basis[idim][1] > 0 else 0))
        if where == HOST:
            dofs = _as_in(dofs, (-1, N))
            E = dofs.shape[0]
            out = _as_out(out, (E, _n_basis(lower) if 0 <= idim < len(basis) and basis[idim][1] > 0 else 0))
        else:
            E = int(n_elements)
        check(lib.ib200_dof_derivative_batch(self.handle, len(basis), basis_specs(basis), int(idim), E, _ptr(dofs), _ptr(out),
                                             where))
        return out

    # -- K0 ----------------------------------------------------------------------------------------
    def geometry(self, geom_basis, rules, geom_dofs, n_phys, want=("coords", "det", "jinv"), where=HOST, n_elements=None,
                 coords=None, det=None, jinv=None):
        """geom_dofs[E][n_phys][Ng] -> coords[E][n_phys][1+d][Q], det[E][Q], jinv[E][Q][d][n_phys]"""
        Ng, Q, d = _n_basis(geom_basis), _n_points(rules), len(geom_basis)
        if _is_dev(geom_dofs, coords, det, jinv):
            geom_dofs = _dev_in(geom_dofs, (-1, n_phys, Ng), "geom_dofs")
            where, n_elements = DEVICE, geom_dofs.shape[0]
            coords = self._dev_out(coords, (n_elements, n_phys, 1 + d, Q)) if "coords" in want else None
            det = self._dev_out(det, (n_elements, Q)) if "det" in want else None
            jinv = self._dev_out(jinv, (n_elements, Q, d, n_phys)) if "jinv" in want else None
        if where == HOST:
            geom_dofs = _as_in(geom_dofs, (-1, n_phys, Ng))
            E = geom_dofs.shape[0]
            coords = _as_out(coords, (E, n_phys, 1 + d, Q), "coords") if "coords" in want else None
            det = _as_out(det, (E, Q), "det") if "det" in want else None
            jinv = _as_out(jinv, (E, Q, d, n_phys), "jinv") if "jinv" in want else None
        else:
            E = int(n_elements)
        check(lib.ib200_geometry_batch(self.handle, d, n_phys, basis_specs(geom_basis), rule_specs(rules), E,
                                       _ptr(geom_dofs), _ptr(coords), _ptr(det), _ptr(jinv), where))
        return coords, det, jinv

    # -- K2 ----------------------------------------------------------------------------------------
    def mass_matrix(self, basis_in, basis_out, rules, n_elements, weight=None, deriv_dim_in=-1, out=None, where=HOST):
        """-> out[E][N_out][N_in]   (ib200_mass_matrix_batch); weight[E][Q] or None"""
        Nin, Nout, Q = _n_basis(basis_in), _n_basis(basis_out), _n_points(rules)
        E = int(n_elements)
        if where == HOST:
            if weight is not None:
                weight = _as_in(weight, (E, Q))
            out = _as_out(out, (E, Nout, Nin))
        check(lib.ib200_mass_matrix_batch(self.handle, len(rules), basis_specs(basis_in), basis_specs(basis_out),
                                          rule_specs(rules), E, _ptr(weight), int(deriv_dim_in), _ptr(out), where))
        return out

    def mass_matrix_mapped(self, basis_in, basis_out, rules, geom_basis, geom_dofs, n_phys, idx_in=-1, idx_out=0, out=None,
                           where=HOST, n_elements=None):
        """geometry fused: geom_dofs[E][n_phys][Ng] -> out[E][N_out][N_in]   (ib200_mass_matrix_mapped_batch)"""
        Nin, Nout, Ng = _n_basis(basis_in), _n_basis(basis_out), _n_basis(geom_basis)
        if _is_dev(geom_dofs, out):
            geom_dofs = _dev_in(geom_dofs, (-1, n_phys, Ng), "geom_dofs")
            where, n_elements = DEVICE, geom_dofs.shape[0]
            out = self._dev_out(out, (n_elements, Nout, Nin))
        if where == HOST:
            geom_dofs = _as_in(geom_dofs, (-1, n_phys, Ng))
            E = geom_dofs.shape[0]
            out = _as_out(out, (E, Nout, Nin))
        else:
            E = int(n_elements)
        check(lib.ib200_mass_matrix_mapped_batch(self.handle, len(rules), basis_specs(basis_in), basis_specs(basis_out),
                                                 rule_specs(rules), n_phys, basis_specs(geom_basis), E, _ptr(geom_dofs),
                                                 int(idx_in), int(idx_out), _ptr(out), where))
        return out

    # -- K3 ----------------------------------------------------------------------------------------
    def l2_dual(self, basis, rules, fvals, det=None, out=None, where=HOST, n_elements=None):
        """fvals[E][Q] (, det[E][Q]) -> out[E][N]   (ib200_l2_dual_batch)"""
        N, Q = _n_basis(basis), _n_points(rules)
        if where == HOST:
            fvals = _as_in(fvals, (-1, Q))
            E = fvals.shape[0]
            if det is not None:
                det = _as_in(det, (E, Q))
            out = _as_out(out, (E, N))
        else:
            E = int(n_elements)
        check(lib.ib200_l2_dual_batch(self.handle, len(rules), basis_specs(basis), rule_specs(rules), E, _ptr(fvals),
                                      _ptr(det), _ptr(out), where))
        return out

    def l2_dual_mapped(self, basis, rules, geom_basis, geom_dofs, n_phys, fvals, out=None, where=HOST, n_elements=None):
        """geometry fused   (ib200_l2_dual_mapped_batch)"""
        N, Q, Ng = _n_basis(basis), _n_points(rules), _n_basis(geom_basis)
        if _is_dev(geom_dofs, fvals, out):
            fvals = _dev_in(fvals, (-1, Q), "fvals")
            where, n_elements = DEVICE, fvals.shape[0]
            geom_dofs = _dev_in(geom_dofs, (n_elements, n_phys, Ng), "geom_dofs")
            out = self._dev_out(out, (n_elements, N))
        if where == HOST:
            fvals = _as_in(fvals, (-1, Q))
            E = fvals.shape[0]
            geom_dofs = _as_in(geom_dofs, (E, n_phys, Ng))
            out = _as_out(out, (E, N))
        else:
            E = int(n_elements)
        check(lib.ib200_l2_dual_mapped_batch(self.handle, len(rules), basis_specs(basis), rule_specs(rules), n_phys,
                                             basis_specs(geom_basis), E, _ptr(geom_dofs), _ptr(fvals), _ptr(out), where))
        return out

    def l2_system_mapped(self, basis, rules, geom_basis, geom_dofs, n_phys, fvals, mass=None, rhs=None, where=HOST,
                         n_elements=None):
        """K2 + K3 in one pass: (mass[E][N][N], rhs[E][N]) of the L2 projection   (ib200_l2_system_mapped_batch)"""
        N, Q, Ng = _n_basis(basis), _n_points(rules), _n_basis(geom_basis)
        if _is_dev(geom_dofs, fvals, mass, rhs):
            fvals = _dev_in(fvals, (-1, Q), "fvals")
            where, n_elements = DEVICE, fvals.shape[0]
            geom_dofs = _dev_in(geom_dofs, (n_elements, n_phys, Ng), "geom_dofs")
            mass = self._dev_out(mass, (n_elements, N, N))
            rhs = self._dev_out(rhs, (n_elements, N))
        if where == HOST:
            fvals = _as_in(fvals, (-1, Q))
            E = fvals.shape[0]
            geom_dofs = _as_in(geom_dofs, (E, n_phys, Ng))
            mass = _as_out(mass, (E, N, N), "mass")
            rhs = _as_out(rhs, (E, N), "rhs")
        else:
            E = int(n_elements)
        check(lib.ib200_l2_system_mapped_batch(self.handle, len(rules), basis_specs(basis), rule_specs(rules), n_phys,
                                               basis_specs(geom_basis), E, _ptr(geom_dofs), _ptr(fvals), _ptr(mass), _ptr(rhs),
                                               where))
        return mass, rhs

    # -- N4 + K2 -----------------------------------------------------------------------------------
    def kform_mass_mapped(self, basis, k, rules, geom_basis, geom_dofs, n_phys, n_dofs, out=None, where=HOST, n_elements=None):
        """The whole k-form element mass matrix M[E][N][N] (N = n_dofs = kform_dof_count): block (a, b) over the component
        spaces = compute_mass_matrix(space_b, space_a, smap); one launch for bilinear quads   (ib200_kform_mass_mapped_batch)"""
        Ng = _n_basis(geom_basis)
        if _is_dev(geom_dofs, out):
            geom_dofs = _dev_in(geom_dofs, (-1, n_phys, Ng), "geom_dofs")
            where, n_elements = DEVICE, geom_dofs.shape[0]
            out = self._dev_out(out, (n_elements, n_dofs, n_dofs))
        if where == HOST:
            geom_dofs = _as_in(geom_dofs, (-1, n_phys, Ng))
            E = geom_dofs.shape[0]
            out = _as_out(out, (E, n_dofs, n_dofs))
        else:
            E = int(n_elements)
        check(lib.ib200_kform_mass_mapped_batch(self.handle, len(rules), basis_specs(basis), int(k), rule_specs(rules), n_phys,
                                                basis_specs(geom_basis), E, _ptr(geom_dofs), _ptr(out), where))
        return out

    # -- N1 ----------------------------------------------------------------------------------------
    def l2_project_mapped(self, basis, rules, geom_basis, geom_dofs, n_phys, fvals, out=None, where=HOST, n_elements=None):
        """projection_l2_primal of a whole batch: geometry + f samples -> primal DoFs u[E][N]; the mass matrices stay in
        device scratch   (ib200_l2_project_mapped_batch)"""
        N, Q, Ng = _n_basis(basis), _n_points(rules), _n_basis(geom_basis)
        if _is_dev(geom_dofs, fvals, out):
            fvals = _dev_in(fvals, (-1, Q), "fvals")
            where, n_elements = DEVICE, fvals.shape[0]
            geom_dofs = _dev_in(geom_dofs, (n_elements, n_phys, Ng), "geom_dofs")
            out = self._dev_out(out, (n_elements, N))
        if where == HOST:
            fvals = _as_in(fvals, (-1, Q))
            E = fvals.shape[0]
            geom_dofs = _as_in(geom_dofs, (E, n_phys, Ng))
            out = _as_out(out, (E, N))
        else:
            E = int(n_elements)
        check(lib.ib200_l2_project_mapped_batch(self.handle, len(rules), basis_specs(basis), rule_specs(rules), n_phys,
                                                basis_specs(geom_basis), E, _ptr(geom_dofs), _ptr(fvals), _ptr(out), where))
        return out

    def spd_solve(self, mass, rhs, where=HOST, n=None, n_elements=None):
        """Solve mass[e] u = rhs[e] for every element (Cholesky on the GPU); returns u[E][N].  Device arrays are solved in
        place: ``rhs`` is overwritten with u (and returned), ``mass`` with its factor."""
        if _is_dev(mass, rhs):
            if not (isinstance(mass, DeviceArray) and isinstance(rhs, DeviceArray)):
                raise TypeError("spd_solve: mass and rhs must both be device arrays")
            where, n_elements, n = DEVICE, rhs.shape[0], rhs.size // max(rhs.shape[0], 1)
        if where == HOST:
            rhs = np.array(rhs, dtype=np.float64, order="C", copy=True)
            E, n = rhs.shape
            mass = np.array(mass, dtype=np.float64, order="C", copy=True).reshape(E, n, n)
        else:
            E = int(n_elements)
        check(lib.ib200_spd_solve_batch(self.handle, int(n), E, _ptr(mass), _ptr(rhs), where))
        return rhs


    # -- norms ---------------------------------------------------------------------------------------
    def dot(self, a, b, out=None, where=HOST, n=None):
        """sum_i a[i]*b[i] with a fixed reduction tree (ib200_dot).  Host arrays -> float; device pointers -> the
        result is left in ``out`` (device pointer to one double); device arrays -> a one-element device array."""
        if _is_dev(a, b, out):
            if not (isinstance(a, DeviceArray) and isinstance(b, DeviceArray)) or a.size != b.size:
                raise ValueError("dot: operands must be device arrays of the same size")
            where, n = DEVICE, a.size
            out = self._dev_out(out, (1,))
        if where == HOST:
            a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1)
            b = np.ascontiguousarray(b, dtype=np.float64).reshape(-1)
            if a.size != b.size:
                raise ValueError("dot: operands differ in size")
            res = np.zeros(1)
            check(lib.ib200_dot(self.handle, a.size, _ptr(a), _ptr(b), _ptr(res), HOST))
            return float(res[0])
        check(lib.ib200_dot(self.handle, int(n), _ptr(a), _ptr(b), _ptr(out), DEVICE))
        return out


# ---- default context (lazy; one per process) -----------------------------------------------------------
_default_ctx: Context | None = None
_default_lock = threading.Lock()


def device_count() -> int:
    n = C.c_int()
    code = lib.ib200_device_count(C.byref(n))
    return n.value if code == 0 else 0


def default_context() -> Context:
    """The process-wide context on device ``$LOCAL_RANK`` (or 0).  Raises when no GPU is usable --
    the batched path has no CPU implementation."""
    global _default_ctx
    with _default_lock:
        if _default_ctx is None:
            import os

            _default_ctx = Context(int(os.environ.get("INTERP_B200_DEVICE", os.environ.get("LOCAL_RANK", "0"))))
        return _default_ctx


# ---- host tables (no GPU needed) -----------------------------------------------------------------------------
def rule_1d(method, order):
    x = np.empty(order + 1)
    w = np.empty(order + 1)
    acc = C.c_uint()
    check(lib.ib200_rule_1d(rule_specs([(method, order)])[0], x.ctypes.data_as(_lib._dp), w.ctypes.data_as(_lib._dp), C.byref(acc)))
    return x, w, acc.value


def lagrange_roots(btype, order):
    r = np.empty(order + 1)
    check(lib.ib200_lagrange_roots(basis_specs([(btype, order)])[0], r.ctypes.data_as(_lib._dp)))
    return r


def basis_table_1d(basis, rule):
    (bt, bo), (rm, ro) = basis, rule
    B = np.empty((bo + 1, ro + 1))
    D = np.empty((bo + 1, ro + 1))
    check(lib.ib200_basis_table_1d(basis_specs([basis])[0], rule_specs([rule])[0], B.ctypes.data_as(_lib._dp),
                                   D.ctypes.data_as(_lib._dp)))
    return B, D


def incidence_matrix(basis):
    """[n][n+1] derivative-DoF operator of a 1-D basis (ib200_incidence_matrix; src/python/incidence.c:38-110)."""
    bt, bo = basis
    out = np.empty((max(bo, 0), bo + 1))
    check(lib.ib200_incidence_matrix(basis_specs([basis])[0], out.ctypes.data_as(C.c_void_p)))
    return out


def pair_moments(basis_out, basis_in, rule, kmax: int = 1, deriv_in: bool = False) -> np.ndarray:
    """out[k][j][i] = sum_q w_q x_q^k Bout_j(x_q) Bin_i(x_q) (Bin' with ``deriv_in``): the 1-D tables of the Kronecker-sum mass
    kernels (ib200_pair_moments)."""
    out = np.empty((kmax + 1, basis_out[1] + 1, basis_in[1] + 1))
    check(lib.ib200_pair_moments(basis_specs([basis_out])[0], basis_specs([basis_in])[0], rule_specs([rule])[0], int(kmax),
                                 int(bool(deriv_in)), out.ctypes.data_as(C.c_void_p)))
    return out


def fastdiag_tables(basis, rule):
    """(V[n][n], lambda[n], Vw[n][m]) of a 1-D basis under a 1-D rule: V^T A V = I, V^T X V = diag(lambda) for the mass
    matrix A and the coordinate-weighted mass matrix X (ib200_fastdiag_tables; the tables of the fast-diagonalisation
    projection, csrc/fastdiag.c)."""
    n, m = basis[1] + 1, rule[1] + 1
    V, lam, Vw = np.empty((n, n)), np.empty(n), np.empty((n, m))
    check(lib.ib200_fastdiag_tables(basis_specs([basis])[0], rule_specs([rule])[0], V.ctypes.data_as(C.c_void_p),
                                    lam.ctypes.data_as(C.c_void_p), Vw.ctypes.data_as(C.c_void_p)))
    return V, lam, Vw


def integration_space_weights(rules):
    W = np.empty([o + 1 for _, o in rules])
    check(lib.ib200_integration_space_weights(len(rules), rule_specs(rules), W.ctypes.data_as(_lib._dp)))
    return W


def integration_space_nodes(rules):
    X = np.empty([len(rules)] + [o + 1 for _, o in rules])
    check(lib.ib200_integration_space_nodes(len(rules), rule_specs(rules), X.ctypes.data_as(_lib._dp)))
    return X


def values_at_integration_nodes(basis, rules):
    out = np.empty([o + 1 for _, o in rules] + [o + 1 for _, o in basis])
    check(lib.ib200_values_at_integration_nodes(len(rules), basis_specs(basis), rule_specs(rules), out.ctypes.data_as(_lib._dp)))
    return out


def basis_at_points(basis, x, derivative=False):
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.empty(x.shape + (basis[1] + 1,))
    args = (None, out.ctypes.data_as(_lib._dp)) if derivative else (out.ctypes.data_as(_lib._dp), None)
    check(lib.ib200_basis_at_points(basis_specs([basis])[0], x.size, x.ctypes.data_as(_lib._dp), *args))
    return out


def function_space_evaluate(basis, xs):
    xs = [np.ascontiguousarray(x, dtype=np.float64) for x in xs]
    shape = xs[0].shape
    out = np.empty(shape + tuple(o + 1 for _, o in basis))
    ptrs = (_lib._dp * len(xs))(*[x.ctypes.data_as(_lib._dp) for x in xs])
    check(lib.ib200_function_space_evaluate(len(basis), basis_specs(basis), xs[0].size, ptrs, out.ctypes.data_as(_lib._dp)))
    return out


# ---- src/polynomials on caller-supplied nodes (host table layer) ----------------------------------------------------
def lagrange_samples(roots, x, deriv: int) -> np.ndarray:
    """out[len(x)][len(roots)] = d^deriv l_j / dx^deriv at x (ib200_lagrange_samples)."""
    roots = np.ascontiguousarray(roots, dtype=np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.empty((x.size, roots.size))
    check(lib.ib200_lagrange_samples(roots.size, roots.ctypes.data_as(C.c_void_p), x.size, x.ctypes.data_as(C.c_void_p),
                                     int(deriv), out.ctypes.data_as(C.c_void_p)))
    return out


def gauss_lobatto(order: int, max_iter: int, tol: float):
    """(nodes, weights, non-converged node count) of the (order+1)-point Gauss-Lobatto rule (ib200_gauss_lobatto)."""
    nodes, weights, bad = np.empty(order + 1), np.empty(order + 1), C.c_int(0)
    check(lib.ib200_gauss_lobatto(int(order), int(max_iter), float(tol), nodes.ctypes.data_as(C.c_void_p),
                                  weights.ctypes.data_as(C.c_void_p), C.byref(bad)))
    return nodes, weights, int(bad.value)


def bernstein_samples(order: int, t) -> np.ndarray:
    t = np.ascontiguousarray(t, dtype=np.float64)
    out = np.empty((t.size, order + 1))
    check(lib.ib200_bernstein_samples(int(order), t.size, t.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
    return out


def bernstein_from_power_series(coeffs: np.ndarray) -> np.ndarray:
    """In place on a contiguous float64 array (ib200_bernstein_from_power_series); returns it."""
    check(lib.ib200_bernstein_from_power_series(coeffs.size, coeffs.ctypes.data_as(C.c_void_p)))
    return coeffs
