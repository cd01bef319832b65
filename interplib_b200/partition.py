"""Element partitioning across the GPUs of one box.

Every element of the path is independent (no inter-element term in A5-A12; the reference has no global
assembly), so the batch is cut into contiguous blocks of the leading element index, one block per rank, one
process per GPU (``torchrun``).  The compute needs NO communication; after it, and only on request,

* ``gather_slabs``  -- all-gather of the per-rank ``[E_g, N]`` DoF / right-hand-side slabs into ``[E, N]``
                       (uneven tails are padded for the collective and stripped afterwards),
* ``reduce_sums``   -- all-reduce(sum) of a handful of FP64 scalars (projection norms),

run over ``torch.distributed`` -- NCCL over NVLink on the GPUs, gloo in the CPU tests.  Mass matrices are never
gathered (they stay sharded; C5 holds 84 GB of them).  torch is plumbing here: device memory for the collective
buffers, streams and the process group; all arithmetic is in ``libinterp_b200.so``.
"""
from __future__ import annotations

import dataclasses

import numpy as np


def element_range(n_elements: int, world: int, rank: int) -> tuple[int, int]:
    """(first, count) of rank ``rank``: GPU g owns ``[g*ceil(E/G), min(E, (g+1)*ceil(E/G)))``."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside a world of {world}")
    if n_elements < 0:
        raise ValueError("n_elements must be non-negative")
    per = -(-n_elements // world)
    first = min(n_elements, rank * per)
    return first, min(n_elements, first + per) - first


def _dist():
    import torch.distributed as dist

    return dist if dist.is_available() and dist.is_initialized() else None


@dataclasses.dataclass
class ElementPartition:
    """Block partition of ``n_elements`` over the ranks of a ``torch.distributed`` group (world 1 when no group
    is initialised: every call then degenerates to the identity)."""

    n_elements: int
    world: int = 1
    rank: int = 0
    group: object = None

    @classmethod
    def from_environment(cls, n_elements: int, group=None) -> "ElementPartition":
        dist = _dist()
        if dist is None:
            return cls(n_elements)
        return cls(n_elements, dist.get_world_size(group), dist.get_rank(group), group)

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

    # -- collectives ------------------------------------------------------------------------------------
    def gather_slabs(self, local):
        """All-gather per-rank slabs ``[count, ...]`` -> ``[n_elements, ...]`` on every rank.

        ``local`` is a torch tensor (CUDA for NCCL, CPU for gloo) or a NumPy array (gloo / world 1); the result
        has the same kind.  Issued on the current stream, no host synchronisation."""
        import torch

        as_numpy = isinstance(local, np.ndarray)
        t = torch.from_numpy(np.ascontiguousarray(local)) if as_numpy else local.contiguous()
        if t.shape[0] != self.count:
            raise ValueError(f"rank {self.rank} owns {self.count} elements, slab has {t.shape[0]}")
        dist = _dist()
        if self.world == 1 or dist is None:
            return local
        tail = tuple(t.shape[1:])
        even = self.max_count * self.world == self.n_elements
        if even:
            out = torch.empty((self.n_elements,) + tail, dtype=t.dtype, device=t.device)
            dist.all_gather_into_tensor(out, t, group=self.group)
        else:
            padded = torch.zeros((self.max_count,) + tail, dtype=t.dtype, device=t.device)
            padded[: self.count] = t
            buf = torch.empty((self.world * self.max_count,) + tail, dtype=t.dtype, device=t.device)
            dist.all_gather_into_tensor(buf, padded, group=self.group)
            buf = buf.view((self.world, self.max_count) + tail)
            out = torch.cat([buf[r, :c] for r, c in enumerate(self.counts)])
        return out.numpy() if as_numpy else out

    def reduce_sums(self, values):
        """All-reduce(sum) of FP64 scalars (projection norms).  ``values``: sequence of floats, NumPy array or torch
        tensor; returns the same kind with the sums over all ranks."""
        import torch

        dist = _dist()
        if isinstance(values, torch.Tensor):
            if self.world > 1 and dist is not None:
                dist.all_reduce(values, op=dist.ReduceOp.SUM, group=self.group)
            return values
        arr = np.atleast_1d(np.asarray(values, dtype=np.float64)).copy()
        if self.world > 1 and dist is not None:
            t = torch.from_numpy(arr)
            backend = dist.get_backend(self.group)
            if backend == "nccl":
                t = t.cuda()
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            arr = t.cpu().numpy()
        return arr


def project_l2_partitioned(ctx, part: ElementPartition, basis, rules, geom_basis, geom_dofs, n_phys, fvals, *, gather=True):
    """L2 projection of sampled values over a partitioned batch, device resident.

    ``geom_dofs`` / ``fvals`` are THIS RANK's slabs as CUDA torch tensors (``[count, n_phys, N_g]``, ``[count, Q]``).
    Per rank: fused-geometry mass matrices (K2), right-hand sides (K3), batched Cholesky solve (N1), and the local
    parts of the norms ``sum_e b_e.u_e`` (= ||Pi f||^2) -- all on the context stream, which is torch's current stream
    so that the NCCL collectives that follow are ordered behind the kernels without a host sync.  Returns
    ``(dofs, norms)``: ``dofs`` = the gathered ``[E, N]`` primal DoFs on every rank (the local slab when
    ``gather=False``), ``norms`` = ``{"proj_sq": sum over all elements of b.u}``.
    """
    import torch

    from ._lib import DEVICE

    E = int(fvals.shape[0])
    if E != part.count or int(geom_dofs.shape[0]) != E:
        raise ValueError("slabs do not match this rank's share of the partition")
    N = int(np.prod([o + 1 for _, o in basis]))
    dev = fvals.device
    ctx.set_stream(torch.cuda.current_stream(dev).cuda_stream)
    try:
        mass = torch.empty((E, N, N), dtype=torch.float64, device=dev)
        rhs = torch.empty((E, N), dtype=torch.float64, device=dev)
        ctx.mass_matrix_mapped(basis, basis, rules, geom_basis, geom_dofs.data_ptr(), n_phys, out=mass.data_ptr(), where=DEVICE,
                               n_elements=E)
        ctx.l2_dual_mapped(basis, rules, geom_basis, geom_dofs.data_ptr(), n_phys, fvals.data_ptr(), out=rhs.data_ptr(),
                           where=DEVICE, n_elements=E)
        dual = rhs.clone()
        ctx.spd_solve(mass.data_ptr(), rhs.data_ptr(), where=DEVICE, n=N, n_elements=E)  # rhs <- primal DoFs
        norm = torch.zeros(1, dtype=torch.float64, device=dev)
        ctx.dot(dual.data_ptr(), rhs.data_ptr(), out=norm.data_ptr(), where=DEVICE, n=E * N)
        del mass
        norms = part.reduce_sums(norm)
        dofs = part.gather_slabs(rhs) if gather else rhs
    finally:
        ctx.set_stream(None)
    return dofs, {"proj_sq": norms}
