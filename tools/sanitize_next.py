#!/usr/bin/env python3
"""Small-case driver over the round-1f kernels (register, two-block and blocked Cholesky, bulk-copy derivative kernel,
register point kernels, fused projection) with ragged sizes: run it plainly for a quick check against NumPy, or under
`compute-sanitizer --tool memcheck|racecheck` on a machine where that tool is available (it is closed on the round's GPU pool)."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import interplib_b200 as ib  # noqa: E402
from interplib_b200 import synthetic as syn  # noqa: E402

ctx = ib.Context(0)
rng = np.random.default_rng(0)
for n, E in ((5, 37), (25, 300), (32, 9), (33, 5), (49, 21), (100, 3), (130, 3), (289, 2)):
    G = rng.standard_normal((E, n, n))
    M = G @ np.swapaxes(G, 1, 2) + n * np.eye(n)
    b = rng.standard_normal((E, n))
    u = ctx.spd_solve(M, b.copy())
    err = np.abs(u - np.linalg.solve(M, b[..., None])[..., 0]).max()
    print("spd_solve", n, E, err)
for basis, E in (([("lagrange-gauss-lobatto", 6)] * 2, 1000), ([("legendre", 3), ("bernstein", 5)], 333), ([("lagrange-gauss", 12)], 77)):
    N = int(np.prod([o + 1 for _, o in basis]))
    dofs = rng.standard_normal((E, N))
    for k in range(len(basis)):
        print("dof_derivative", basis, k, ctx.dof_derivative(basis, dofs, k).shape)
for orders, E, P in (((6, 6), 131, 64), ((3, 4), 50, 33), ((7, 7), 9, 200)):
    basis = [("lagrange-gauss-lobatto", orders[0]), ("lagrange-gauss", orders[1])]
    N = (orders[0] + 1) * (orders[1] + 1)
    out = ctx.reconstruct_points(basis, rng.standard_normal((E, N)), rng.uniform(-1, 1, (E, 2, P)), [0, 1, 2, 3])
    print("reconstruct_points", orders, out.shape)
cfg = syn.CONFIGS["c5"]
E = 3000
geom = syn.geometry(cfg, first=0, count=E)
u = ctx.l2_project_mapped(list(cfg.basis), list(cfg.rules), list(cfg.geom_basis), geom, 2, rng.standard_normal((E, cfg.Q)))
print("l2_project", u.shape, np.isfinite(u).all())
ctx.close()
