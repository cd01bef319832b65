#!/usr/bin/env python3
"""Launches each round-2 kernel a few times on device-resident data (for `ncu -k regex:<kernel>` captures):
    python tools/prof_new_kernels.py kron3 | gradmirror | fd_c4"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import interplib_b200 as ib  # noqa: E402
from interplib_b200 import DEVICE, synthetic as syn  # noqa: E402

GLL, GAUSS, LU = "lagrange-gauss-lobatto", "gauss", "lagrange-uniform"
what = sys.argv[1]
ctx = ib.Context(0)
rng = np.random.default_rng(1)
if what == "kron3":
    p = 3
    basis, rules, gb = [(GLL, p)] * 3, [(GAUSS, p + 1)] * 3, [(LU, 1)] * 3
    N = (p + 1) ** 3
    E = int(8e9 / (8 * N * N)) // 256 * 256
    corners = np.array([[[2 * i0 - 1, 2 * i1 - 1, 2 * i2 - 1] for i0 in range(2) for i1 in range(2) for i2 in range(2)]], dtype=float)
    g = np.ascontiguousarray((corners + rng.uniform(-0.2, 0.2, (E, 8, 3))).transpose(0, 2, 1))
    d_g, d_M = ctx.alloc(g.nbytes).upload(g), ctx.alloc(E * N * N * 8)
    run = lambda: ctx.mass_matrix_mapped(basis, basis, rules, gb, d_g.ptr, 3, out=d_M.ptr, where=DEVICE, n_elements=E)  # noqa: E731
elif what == "gradmirror":
    p, E = 16, 1 << 14
    basis, rules, gb = [(GLL, p)] * 2, [(GAUSS, p + 1)] * 2, [(LU, 1)] * 2
    N = (p + 1) ** 2
    geom = np.ascontiguousarray(syn.quad_mesh_geometry(128 * 128, seed=1).reshape(-1, 2, 4)[:E])
    d_g, d_M = ctx.alloc(geom.nbytes).upload(geom), ctx.alloc(E * N * N * 8)
    ctx.set_kron_gradient(True)
    run = lambda: ctx.mass_matrix_mapped(basis, basis, rules, gb, d_g.ptr, 2, idx_in=1, idx_out=0, out=d_M.ptr, where=DEVICE, n_elements=E)  # noqa: E731
else:
    p, E = 16, 1 << 16
    basis, rules, gb = [(GLL, p)] * 2, [(GAUSS, p + 1)] * 2, [(LU, 1)] * 2
    N, Q = (p + 1) ** 2, (p + 2) ** 2
    geom = np.ascontiguousarray(syn.quad_mesh_geometry(256 * 256, seed=1).reshape(-1, 2, 4)[:E])
    f = rng.standard_normal((E, Q))
    d_g, d_f, d_u = ctx.alloc(geom.nbytes).upload(geom), ctx.alloc(f.nbytes).upload(f), ctx.alloc(E * N * 8)
    run = lambda: ctx.l2_project_mapped(basis, rules, gb, d_g.ptr, 2, d_f.ptr, out=d_u.ptr, where=DEVICE, n_elements=E)  # noqa: E731
for _ in range(4):
    run()
ctx.synchronize()
