"""Multi-GPU host logic on CPU: the element partition (bit-exact integer bookkeeping) and the two collectives of
SURVEY 8e -- all-gather of DoF slabs, all-reduce of norms -- over gloo with world size 2 (and 3, uneven tails).  The
product path runs the same bookkeeping over NCCL through the C ABI (tests/test_partition_gpu.py); gloo stands in for it
here through tests/_gloo_comm.py."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from interplib_b200.partition import ElementPartition, element_range  # noqa: E402


@pytest.mark.parametrize("E", [0, 1, 7, 8, 1000, 1 << 24])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_block_partition_covers_every_element_once(E, world):
    ranges = [element_range(E, world, r) for r in range(world)]
    per = -(-E // world)
    pos = 0
    for first, count in ranges:
        assert 0 <= count <= per and (first == pos or count == 0)
        pos += count
    assert pos == E
    part = ElementPartition(E, world, 0)
    assert part.counts == [c for _, c in ranges] and part.offsets == [f for f, _ in ranges]
    if E:
        for e in {0, E // 2, E - 1}:
            r = part.owner(e)
            assert ranges[r][0] <= e < ranges[r][0] + ranges[r][1]
    with pytest.raises(ValueError):
        element_range(E, world, world)


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, E: int, N: int, tmp: str):
    import torch.distributed as dist

    from _gloo_comm import GlooComm

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        part = ElementPartition.from_comm(E, GlooComm())
        assert (part.world, part.rank) == (world, rank)
        full = np.random.default_rng(42).random((E, N))  # same on every rank
        mine = part.local(full)
        assert mine.shape == (part.count, N)
        # each rank "computes" only its slab; the gather must reproduce the full array bit for bit, in element order
        gathered = part.gather_slabs(np.ascontiguousarray(mine))
        assert gathered.shape == (E, N) and np.array_equal(gathered, full)
        # pipelined form: sub-slabs of 3 elements of every rank's slab land at their final places
        piece = np.zeros((E, N))
        for lo in range(0, part.max_count, 3):
            counts, offsets = part.slab_layout(N, lo, 3)
            part.comm.allgatherv(np.ascontiguousarray(mine[lo : lo + 3]), piece, counts, offsets)
        assert np.array_equal(piece, full)
        # norms: per-rank partial sums -> all-reduce
        sums = part.reduce_sums([float(np.sum(mine * mine)), float(part.count)])
        assert sums[1] == E and abs(sums[0] - np.sum(full * full)) <= 1e-12 * np.sum(full * full)
        with pytest.raises(ValueError):
            part.gather_slabs(np.zeros((part.count + 1, N)))
        Path(tmp, f"ok{rank}").write_text("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,E", [(2, 1000), (2, 1001), (3, 10), (2, 1)])
def test_gather_and_reduce_over_gloo(tmp_path, world, E):
    import torch.multiprocessing as mp

    mp.spawn(_worker, args=(world, _free_port(), E, 25, str(tmp_path)), nprocs=world, join=True)
    assert sorted(p.name for p in tmp_path.iterdir()) == [f"ok{r}" for r in range(world)]


def test_single_process_partition_is_the_identity():
    part = ElementPartition.from_comm(5)
    a = np.arange(10.0).reshape(5, 2)
    assert part.world == 1 and part.local(a) is not None and np.array_equal(part.gather_slabs(a), a)
    assert part.reduce_sums([1.5, 2.5]).tolist() == [1.5, 2.5]
