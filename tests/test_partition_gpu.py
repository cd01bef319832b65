"""Multi-GPU path on real devices: element-partitioned L2 projection with one process per GPU, NCCL
all-gather of the primal DoF slabs and all-reduce of the projection norm.  Needs >= 2 GPUs (skipped on a 1-GPU box;
the same host logic runs on gloo in tests/test_partition_cpu.py).  Checked against the CPU oracle on the full batch."""
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
    import torch
    import torch.distributed as dist

    import interplib_b200 as ib
    from interplib_b200 import synthetic as syn
    from interplib_b200.partition import ElementPartition, project_l2_partitioned

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        cfg = syn.CONFIGS["c5"]
        basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
        part = ElementPartition.from_environment(E)
        geom = syn.geometry(cfg, first=123, count=E)          # same on every rank
        f = np.random.default_rng(5).standard_normal((E, cfg.Q))
        ctx = ib.Context(rank)
        g_loc = torch.from_numpy(np.ascontiguousarray(part.local(geom))).cuda()
        f_loc = torch.from_numpy(np.ascontiguousarray(part.local(f))).cuda()
        dofs, norms = project_l2_partitioned(ctx, part, basis, rules, gb, g_loc, cfg.n_phys, f_loc)
        torch.cuda.synchronize()
        assert tuple(dofs.shape) == (E, cfg.N)
        np.save(Path(tmp, f"dofs{rank}.npy"), dofs.cpu().numpy())
        np.save(Path(tmp, f"norm{rank}.npy"), norms["proj_sq"].cpu().numpy())
        ctx.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("E", [4096, 4097])
def test_partitioned_projection_over_nccl(tmp_path, orc, E):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp

    from interplib_b200 import synthetic as syn

    world = 2
    mp.spawn(_worker, args=(world, _free_port(), E, str(tmp_path)), nprocs=world, join=True)
    cfg = syn.CONFIGS["c5"]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    geom = syn.geometry(cfg, first=123, count=E)
    f = np.random.default_rng(5).standard_normal((E, cfg.Q))
    _, det, _ = orc.geometry(gb, rules, geom, cfg.n_phys, want_coords=False, want_jinv=False)
    M = orc.mass(basis, basis, rules, E, det=det)
    b = orc.l2_dual(basis, rules, f, det=det)
    u = np.linalg.solve(M, b[..., None])[..., 0]
    d0, d1 = (np.load(tmp_path / f"dofs{r}.npy") for r in range(world))
    assert np.array_equal(d0, d1)                       # every rank holds the same gathered array, in element order
    scale = np.abs(u).max(axis=1, keepdims=True)
    assert (np.abs(d0 - u) / scale).max() < 1e-11      # cond(M) ~ 1e2 on top of the 1e-12 kernel tolerance
    n0, n1 = (np.load(tmp_path / f"norm{r}.npy") for r in range(world))
    ref = float(np.sum(b * u))
    assert n0[0] == n1[0] and abs(n0[0] - ref) <= 1e-11 * abs(ref)
