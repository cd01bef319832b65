"""Multi-GPU path on real devices: element-partitioned L2 projection with one process per GPU, NCCL
all-gather of the primal DoF slabs and all-reduce of the projection norm -- through the C ABI only (ib200_comm_*,
ib200_allgatherv, ib200_allreduce_sums); no torch anywhere on this path.  The world-size-1 cases run on a 1-GPU box
(NCCL is still what moves the data); the two-rank cases need >= 2 GPUs (`gpurun --gpus 2`).  Checked against the CPU
oracle on the full batch."""
import multiprocessing as mp
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, E: int, tmp: str):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import interplib_b200 as ib
    from interplib_b200 import synthetic as syn
    from interplib_b200.partition import ElementPartition, NcclComm, project_l2_partitioned

    assert "torch" not in sys.modules, "the partitioned product path must not need torch"
    cfg = syn.CONFIGS["c5"]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    ctx = ib.Context(rank)
    comm = NcclComm.from_environment(ctx, port_offset=0)
    info = comm.info()
    assert (info["n_ranks"], info["rank"]) == (world, rank) and info["nccl_version"] >= 22000
    part = ElementPartition.from_comm(E, comm)
    geom = syn.geometry(cfg, first=123, count=E)  # same on every rank
    f = np.random.default_rng(5).standard_normal((E, cfg.Q))
    dofs, norms = project_l2_partitioned(ctx, part, basis, rules, gb, part.local(geom), cfg.n_phys, part.local(f))
    assert dofs.shape == (E, cfg.N)
    # host-array form of the gather (staged through the device) and of the reduction
    again = part.gather_slabs(np.ascontiguousarray(part.local(dofs)))
    assert np.array_equal(again, dofs)
    sums = part.reduce_sums([float(part.count), 1.0])
    assert sums.tolist() == [float(E), float(world)]
    # pipelined form: sub-slabs gathered on the communicator's side stream, joined at the end
    d_all = ctx.alloc(E * cfg.N * 8)
    ctx_local = ctx.alloc(max(part.count, 1) * cfg.N * 8).upload(np.ascontiguousarray(part.local(dofs)))
    step = max(1, part.max_count // 3)
    for lo in range(0, part.max_count, step):
        counts, offsets = part.slab_layout(cfg.N, lo, step)
        comm.allgatherv(int(ctx_local) + min(lo, part.count) * cfg.N * 8, d_all, counts, offsets, overlap=True)
    comm.join()
    ctx.synchronize()
    assert np.array_equal(d_all.download((E, cfg.N)), dofs)
    # registered destination: the same pipelined gather as copy-engine pushes into IPC-mapped peer buffers (NVLink), plus a
    # whole-slab call on the compute stream; must deliver the same bytes as NCCL did
    d_p2p = ctx.alloc(E * cfg.N * 8)
    p2p = comm.register_gather(d_p2p, E * cfg.N * 8)
    assert p2p or world == 1 or os.environ.get("IB200_NO_P2P_GATHER"), "CUDA IPC peer mapping failed on a multi-GPU box"
    for lo in range(0, part.max_count, step):
        counts, offsets = part.slab_layout(cfg.N, lo, step)
        comm.allgatherv(int(ctx_local) + min(lo, part.count) * cfg.N * 8, d_p2p, counts, offsets, overlap=True)
    comm.join()
    ctx.synchronize()
    assert np.array_equal(d_p2p.download((E, cfg.N)), dofs)
    d_p2p.upload(np.zeros((E, cfg.N)))
    comm.allgather_slabs(ctx_local, E, cfg.N, d_p2p)
    ctx.synchronize()
    assert np.array_equal(d_p2p.download((E, cfg.N)), dofs)
    comm.unregister_gather()
    np.save(Path(tmp, f"p2p{rank}.npy"), np.array([1.0 if p2p else 0.0]))
    comm.check()
    np.save(Path(tmp, f"dofs{rank}.npy"), dofs)
    np.save(Path(tmp, f"norm{rank}.npy"), np.array([norms["proj_sq"]]))
    comm.close()
    ctx.close()


def _run(world, E, tmp_path):
    mpc = mp.get_context("spawn")
    port = _free_port()
    procs = [mpc.Process(target=_worker, args=(r, world, port, E, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
    assert [p.exitcode for p in procs] == [0] * world


def _check(world, E, tmp_path, orc):
    from interplib_b200 import synthetic as syn

    cfg = syn.CONFIGS["c5"]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    geom = syn.geometry(cfg, first=123, count=E)
    f = np.random.default_rng(5).standard_normal((E, cfg.Q))
    _, det, _ = orc.geometry(gb, rules, geom, cfg.n_phys, want_coords=False, want_jinv=False)
    M = orc.mass(basis, basis, rules, E, det=det)
    b = orc.l2_dual(basis, rules, f, det=det)
    u = np.linalg.solve(M, b[..., None])[..., 0]
    d = [np.load(tmp_path / f"dofs{r}.npy") for r in range(world)]
    for other in d[1:]:
        assert np.array_equal(d[0], other)  # every rank holds the same gathered array, in element order
    scale = np.abs(u).max(axis=1, keepdims=True)
    assert (np.abs(d[0] - u) / scale).max() < 1e-11  # cond(M) ~ 1e2 on top of the 1e-12 kernel tolerance
    n = [np.load(tmp_path / f"norm{r}.npy")[0] for r in range(world)]
    ref = float(np.sum(b * u))
    assert all(x == n[0] for x in n) and abs(n[0] - ref) <= 1e-11 * abs(ref)


@pytest.mark.parametrize("E", [1000, 1])
def test_world_of_one_still_goes_through_nccl(tmp_path, orc, E):
    _run(1, E, tmp_path)
    _check(1, E, tmp_path, orc)


@pytest.mark.parametrize("E", [4096, 4097])
def test_partitioned_projection_over_nccl(tmp_path, orc, E):
    import interplib_b200 as ib

    if ib.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    _run(2, E, tmp_path)
    _check(2, E, tmp_path, orc)
