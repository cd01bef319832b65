"""Element partitioning across the GPUs of one box.

Every element of the path is independent (no inter-element term in A5-A12; the reference has no global assembly), so
the batch is cut into contiguous blocks of the leading element index, one block per rank, one process (or one
context) per GPU.  The compute needs NO communication; after it, and only on request,

* ``gather_slabs``  -- all-gather of the per-rank ``[E_g, N]`` DoF / right-hand-side slabs into ``[E, N]``
                       (ragged tails need no padding),
* ``reduce_sums``   -- all-reduce(sum) of a handful of FP64 scalars (projection norms),

run through the C ABI (``ib200_allgatherv`` / ``ib200_allreduce_sums``: NCCL over NVLink, bound inside
``libinterp_b200.so``).  Mass matrices are never gathered (they stay sharded; C5 holds 84 GB of them).  Nothing here
imports torch: the communicator is bootstrapped from the 128-byte NCCL id, which rank 0 hands to the others over a
TCP socket (``MASTER_ADDR`` / ``MASTER_PORT`` + 1, the variables ``torchrun`` exports) or by any means the caller
prefers (``NcclComm.create``).

``ElementPartition`` only needs an object with ``n_ranks``, ``rank``, ``allgatherv`` and ``allreduce_sums``; the CPU
tests drive the same bookkeeping with a gloo-backed stand-in (``tests/_gloo_comm.py``).
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import os
import socket
import time

import numpy as np

from . import _lib
from ._lib import check, lib


def element_range(n_elements: int, world: int, rank: int) -> tuple[int, int]:
    """(first, count) of rank ``rank``: GPU g owns ``[g*ceil(E/G), min(E, (g+1)*ceil(E/G)))``
    (``ib200_partition_range``)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside a world of {world}")
    if n_elements < 0:
        raise ValueError("n_elements must be non-negative")
    first, count = C.c_size_t(), C.c_size_t()
    check(lib.ib200_partition_range(int(n_elements), int(world), int(rank), C.byref(first), C.byref(count)))
    return int(first.value), int(count.value)


# ------------------------------------------------------------------------------------------------ communicators
class LocalComm:
    """World of one: every collective is the identity."""

    n_ranks, rank = 1, 0

    def allgatherv(self, send, out, counts, offsets, overlap=False):
        if send is not out:
            np.copyto(np.asarray(out).reshape(-1)[offsets[0] : offsets[0] + counts[0]], np.asarray(send).reshape(-1))

    def allreduce_sums(self, values):
        return values

    def join(self):
        pass


def _exchange_id(rank: int, world: int, make_id, addr: str, port: int, timeout: float = 120.0) -> bytes:
    """Rank 0 serves the NCCL id to the other ranks over TCP."""
    if world == 1:
        return make_id()
    if rank == 0:
        uid = make_id()
        with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as srv:
            srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            srv.bind((addr, port))
            srv.listen(world)
            srv.settimeout(timeout)
            for _ in range(world - 1):
                conn, _ = srv.accept()
                with conn:
                    conn.sendall(uid)
        return uid
    deadline = time.monotonic() + timeout
    while True:
        try:
            with socket.create_connection((addr, port), timeout=5.0) as s:
                buf = b""
                while len(buf) < _lib.COMM_ID_BYTES:
                    part = s.recv(_lib.COMM_ID_BYTES - len(buf))
                    if not part:
                        break
                    buf += part
            if len(buf) == _lib.COMM_ID_BYTES:
                return buf
        except OSError:
            pass
        if time.monotonic() > deadline:
            raise TimeoutError(f"rank {rank}: no NCCL id from rank 0 at {addr}:{port}")
        time.sleep(0.05)


class NcclComm:
    """One NCCL communicator bound to a :class:`~interplib_b200.batch.Context` (``ib200_comm_create``).  All collectives
    take device pointers (``DeviceBuffer`` or int) and are queued on the context stream."""

    def __init__(self, ctx, n_ranks: int, rank: int, uid: bytes):
        if len(uid) != _lib.COMM_ID_BYTES:
            raise ValueError(f"the NCCL id has {_lib.COMM_ID_BYTES} bytes, got {len(uid)}")
        h = C.c_void_p()
        buf = C.create_string_buffer(bytes(uid), _lib.COMM_ID_BYTES)
        check(lib.ib200_comm_create(ctx.handle, int(n_ranks), int(rank), buf, C.byref(h)))
        self.handle, self.ctx, self.n_ranks, self.rank = h, ctx, int(n_ranks), int(rank)

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(_lib.COMM_ID_BYTES)
        check(lib.ib200_comm_unique_id(buf))
        return buf.raw

    @classmethod
    def from_environment(cls, ctx, port_offset: int = 1) -> "NcclComm":
        """Rank / world size / rendezvous address from the variables ``torchrun`` sets (RANK, WORLD_SIZE, MASTER_ADDR,
        MASTER_PORT); the id travels over ``MASTER_PORT + port_offset`` ($IB200_COMM_PORT overrides)."""
        rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
        addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
        port = int(os.environ.get("IB200_COMM_PORT", int(os.environ.get("MASTER_PORT", "29500")) + port_offset))
        return cls(ctx, world, rank, _exchange_id(rank, world, cls.unique_id, addr, port))

    def info(self) -> dict:
        n, r, v = C.c_int(), C.c_int(), C.c_int()
        check(lib.ib200_comm_info(self.handle, C.byref(n), C.byref(r), C.byref(v)))
        return {"n_ranks": n.value, "rank": r.value, "nccl_version": v.value}

    def check(self) -> None:
        """Raises when NCCL reports an asynchronous failure (``ncclCommGetAsyncError``)."""
        check(lib.ib200_comm_check(self.handle))

    def allgatherv(self, send, out, counts, offsets, overlap: bool = False) -> None:
        """counts / offsets in doubles, one entry per rank; ``send`` / ``out`` device pointers."""
        n = self.n_ranks
        cnt = (C.c_size_t * n)(*[int(c) for c in counts])
        off = (C.c_size_t * n)(*[int(o) for o in offsets])
        check(lib.ib200_allgatherv(self.handle, C.c_void_p(int(send) if send is not None else 0), C.c_void_p(int(out)), cnt, off,
                                   int(bool(overlap))))

    def allgather_slabs(self, local, n_elements: int, elem_doubles: int, out) -> None:
        check(lib.ib200_allgather_slabs(self.handle, C.c_void_p(int(local) if local is not None else 0), int(n_elements),
                                        int(elem_doubles), C.c_void_p(int(out))))

    def register_gather(self, out, nbytes: int) -> bool:
        """Collective: map every rank's gather destination ``out`` (base of a device allocation of ``nbytes``) into every
        other rank (CUDA IPC); all-gathers into ``out`` then go as copy-engine pushes over NVLink instead of NCCL kernels.
        False when peer mapping is unavailable (the NCCL path stays in use)."""
        res = lib.ib200_comm_register_gather(self.handle, C.c_void_p(int(out)), int(nbytes))
        if res == 103:  # IB200_ERROR_UNSUPPORTED
            return False
        check(res)
        return True

    def unregister_gather(self) -> None:
        check(lib.ib200_comm_unregister_gather(self.handle))

    def join(self) -> None:
        """Make the context stream wait for the collectives that were sent to the side stream (``overlap=True``)."""
        check(lib.ib200_comm_join(self.handle))

    def allreduce_sums(self, values, n: int | None = None):
        """In-place sum over ranks.  Device pointer + ``n``, or a host sequence (staged through the device)."""
        if isinstance(values, (int, np.integer)) or hasattr(values, "ptr"):
            check(lib.ib200_allreduce_sums(self.handle, C.c_void_p(int(values)), int(n)))
            return values
        arr = np.atleast_1d(np.asarray(values, dtype=np.float64)).copy()
        buf = self.ctx.alloc(arr.nbytes).upload(arr)
        check(lib.ib200_allreduce_sums(self.handle, C.c_void_p(int(buf)), arr.size))
        res = buf.download(arr.shape)
        buf.free()
        return res

    def close(self) -> None:
        if self.handle:
            check(lib.ib200_comm_destroy(self.handle))
            self.handle = None


# ------------------------------------------------------------------------------------------------ the partition
@dataclasses.dataclass
class ElementPartition:
    """Block partition of ``n_elements`` over the ranks of a communicator (world 1 without one: every call then
    degenerates to the identity)."""

    n_elements: int
    world: int = 1
    rank: int = 0
    comm: object = None

    @classmethod
    def from_comm(cls, n_elements: int, comm=None) -> "ElementPartition":
        if comm is None:
            return cls(n_elements)
        return cls(n_elements, comm.n_ranks, comm.rank, comm)

    # -- index bookkeeping (bit-exact: pure integers) ------------------------------------------------
    @property
    def counts(self) -> list[int]:
        return [element_range(self.n_elements, self.world, r)[1] for r in range(self.world)]

    @property
    def offsets(self) -> list[int]:
        return [element_range(self.n_elements, self.world, r)[0] for r in range(self.world)]

    @property
    def first(self) -> int:
        return element_range(self.n_elements, self.world, self.rank)[0]

    @property
    def count(self) -> int:
        return element_range(self.n_elements, self.world, self.rank)[1]

    @property
    def max_count(self) -> int:
        return -(-self.n_elements // self.world)

    def local(self, array):
        """This rank's slab ``array[first : first + count]`` of a full ``[E, ...]`` array."""
        if len(array) != self.n_elements:
            raise ValueError(f"expected {self.n_elements} elements, got {len(array)}")
        return array[self.first : self.first + self.count]

    def owner(self, element: int) -> int:
        if not 0 <= element < self.n_elements:
            raise IndexError(element)
        return element // self.max_count

    def slab_layout(self, elem_doubles: int, sub_first: int = 0, sub_count: int | None = None):
        """(counts, offsets) in doubles of an all-gather where every rank contributes the elements
        ``[sub_first, sub_first + sub_count)`` OF ITS OWN SLAB (clipped to the slab) -- the whole slab by default.
        Sub-slabs are what a pipelined projection gathers while the next sub-slab is being computed."""
        counts, offsets = [], []
        for first, cnt in zip(self.offsets, self.counts):
            lo = min(cnt, sub_first)
            hi = cnt if sub_count is None else min(cnt, sub_first + sub_count)
            counts.append((hi - lo) * elem_doubles)
            offsets.append((first + lo) * elem_doubles)
        return counts, offsets

    # -- collectives ------------------------------------------------------------------------------------
    def gather_slabs(self, local, elem_doubles: int | None = None, out=None):
        """All-gather per-rank slabs ``[count, ...]`` -> ``[n_elements, ...]`` on every rank.

        ``local`` is a NumPy array (staged through the device for an :class:`NcclComm`, handed over as it is to a host
        communicator) or a device pointer / ``DeviceBuffer`` with ``elem_doubles`` given; the result has the same kind
        (``out`` = device buffer to fill, allocated when missing).  Device calls are queued on the context stream."""
        if isinstance(local, np.ndarray):
            if local.shape[0] != self.count:
                raise ValueError(f"rank {self.rank} owns {self.count} elements, slab has {local.shape[0]}")
            if self.world == 1 or self.comm is None:
                return local
            tail = tuple(local.shape[1:])
            per = int(np.prod(tail, dtype=np.int64))
            counts, offsets = self.slab_layout(per)
            src = np.ascontiguousarray(local, dtype=np.float64)
            if isinstance(self.comm, NcclComm):
                ctx = self.comm.ctx
                d_out = ctx.alloc(max(self.n_elements * per * 8, 8))
                d_loc = ctx.alloc(max(src.nbytes, 8)).upload(src)
                self.comm.allgatherv(d_loc, d_out, counts, offsets)
                res = d_out.download((self.n_elements,) + tail)
                d_loc.free()
                d_out.free()
                return res
            res = np.empty((self.n_elements,) + tail)
            self.comm.allgatherv(src, res, counts, offsets)
            return res
        if elem_doubles is None:
            raise ValueError("device slabs need elem_doubles")
        if self.world == 1 or self.comm is None or isinstance(self.comm, LocalComm):
            return local if out is None else out
        if out is None:
            out = self.comm.ctx.alloc(max(self.n_elements * elem_doubles * 8, 8))
        counts, offsets = self.slab_layout(elem_doubles)
        self.comm.allgatherv(local, out, counts, offsets)
        return out

    def reduce_sums(self, values, n: int | None = None):
        """All-reduce(sum) of FP64 scalars (projection norms): a host sequence -> NumPy array of the sums, or a device
        pointer with ``n`` -> reduced in place."""
        if self.world == 1 or self.comm is None:
            return values if n is not None else np.atleast_1d(np.asarray(values, dtype=np.float64)).copy()
        if n is not None:
            return self.comm.allreduce_sums(values, n)
        return np.asarray(self.comm.allreduce_sums(np.atleast_1d(np.asarray(values, dtype=np.float64)).copy()))


def project_l2_partitioned(ctx, part: ElementPartition, basis, rules, geom_basis, geom_dofs, n_phys, fvals, *, gather=True):
    """L2 projection of sampled values over a partitioned batch (``projection_l2_primal`` of every element,
    python/interplib/integration.py:171-223, element-partitioned per SURVEY 8e).

    ``geom_dofs`` / ``fvals`` are THIS RANK's slabs as NumPy arrays (``[count, n_phys, N_g]``, ``[count, Q]``); they are
    uploaded once and everything then stays on the device: right-hand sides (K3), the fused assembly + solve (K2 + N1),
    the local part of ``sum_e b_e.u_e`` (= ||Pi f||^2), the NCCL all-reduce of that norm and the all-gather of the
    primal DoFs -- all queued on the context stream, one synchronisation at the end.  Returns ``(dofs, norms)``:
    ``dofs`` = the gathered ``[E, N]`` primal DoFs (the local ``[count, N]`` slab with ``gather=False``),
    ``norms = {"proj_sq": float}``."""
    from ._lib import DEVICE

    fvals = np.ascontiguousarray(fvals, dtype=np.float64)
    geom_dofs = np.ascontiguousarray(geom_dofs, dtype=np.float64)
    E = int(fvals.shape[0])
    if E != part.count or int(geom_dofs.shape[0]) != E:
        raise ValueError("slabs do not match this rank's share of the partition")
    N = int(np.prod([o + 1 for _, o in basis]))
    d_g = ctx.alloc(max(geom_dofs.nbytes, 8)).upload(geom_dofs)
    d_f = ctx.alloc(max(fvals.nbytes, 8)).upload(fvals)
    d_b = ctx.alloc(max(E * N * 8, 8))
    d_all = ctx.alloc(max(part.n_elements * N * 8, 8)) if gather else None
    # the primal DoFs are written straight into this rank's place of the gathered array (in-place all-gather)
    d_u_ptr = int(d_all) + part.first * N * 8 if gather else None
    d_u = None if gather else ctx.alloc(max(E * N * 8, 8))
    u_ptr = d_u_ptr if gather else int(d_u)
    d_norm = ctx.alloc(8)
    try:
        ctx.l2_dual_mapped(basis, rules, geom_basis, int(d_g), n_phys, int(d_f), out=int(d_b), where=DEVICE, n_elements=E)
        ctx.l2_project_mapped(basis, rules, geom_basis, int(d_g), n_phys, int(d_f), out=u_ptr, where=DEVICE, n_elements=E)
        ctx.dot(int(d_b), u_ptr, out=int(d_norm), where=DEVICE, n=E * N)
        part.reduce_sums(int(d_norm), 1)
        if gather:
            part.gather_slabs(u_ptr, N, out=d_all)
        ctx.synchronize()
        norm = float(d_norm.download((1,))[0])
        dofs = d_all.download((part.n_elements, N)) if gather else d_u.download((E, N))
    finally:
        for buf in (d_g, d_f, d_b, d_all, d_u, d_norm):
            if buf is not None:
                buf.free()
    return dofs, {"proj_sq": norm}
