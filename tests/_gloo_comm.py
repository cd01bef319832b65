"""Test stand-in for the NCCL communicator: the same four members (`n_ranks`, `rank`, `allgatherv`, `allreduce_sums`)
over torch.distributed's gloo backend on host arrays, so that the element-partition bookkeeping of
interplib_b200.partition (counts, offsets, ragged tails, sub-slab layouts) runs with world size > 1 on a CPU-only
machine.  Test infrastructure only -- the product path talks to NCCL through the C ABI."""
import numpy as np
import torch
import torch.distributed as dist


class GlooComm:
    def __init__(self):
        self.n_ranks, self.rank = dist.get_world_size(), dist.get_rank()

    def allgatherv(self, send, out, counts, offsets, overlap=False):
        flat = np.asarray(out).reshape(-1)
        mine = torch.from_numpy(np.ascontiguousarray(send, dtype=np.float64).reshape(-1))
        assert mine.numel() == counts[self.rank]
        for r in range(self.n_ranks):  # one broadcast per rank, like the ragged NCCL path of ib200_allgatherv
            buf = mine.clone() if r == self.rank else torch.empty(counts[r], dtype=torch.float64)
            if counts[r]:
                dist.broadcast(buf, src=r)
            flat[offsets[r] : offsets[r] + counts[r]] = buf.numpy()

    def allreduce_sums(self, values):
        t = torch.from_numpy(np.atleast_1d(np.asarray(values, dtype=np.float64)).copy())
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.numpy()

    def join(self):
        pass
