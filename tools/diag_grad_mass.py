import sys; sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
import interplib_b200 as ib
from interplib_b200 import synthetic as syn
from oracle import oracle as orc
from conftest import rel_err
import test_parity_gpu as T
orc.build()
ctx = ib.Context(0)
rng = np.random.default_rng(13)
for bi, bo, rules in T.MASS_CASES:
    d = len(rules)
    E = 67 if np.prod([o + 1 for _, o in bi]) * np.prod([o + 1 for _, o in bo]) < 20000 else 11
    gb = [("lagrange-uniform", 1)] * d if d > 1 else [("bernstein", 1)]
    base = syn.quad_mesh_geometry(64, seed=3).reshape(64, 2, 4) if d == 2 else None
    if d == 2:
        g = np.concatenate([base, base[:3] * 0.5])[:E]
    elif d == 1:
        g = np.sort(rng.random((E, 1, 2)), axis=2) + np.array([0.0, 0.5])
    else:
        corners = np.array([[[2 * i0 - 1, 2 * i1 - 1, 2 * i2 - 1] for i0 in range(2) for i1 in range(2) for i2 in range(2)]], dtype=float)
        g = (corners + rng.uniform(-0.2, 0.2, (E, 8, 3))).transpose(0, 2, 1).copy()
    _, det, jinv = orc.geometry(gb, rules, g, d)
    kap = np.linalg.cond(jinv.reshape(-1, d, d)).max()
    worst = 0
    for ii in range(d):
        for io in range(d):
            refg = orc.mass(bi, bo, rules, E, det=det, jinv=jinv, n_phys=d, deriv_in=ii, idx_out=io)
            fg = ctx.mass_matrix_mapped(bi, bo, rules, gb, g, d, idx_in=ii, idx_out=io)
            worst = max(worst, rel_err(fg.reshape(E, -1), refg.reshape(E, -1)))
    print(d, bi[0], "worst grad-mass rel err %.2e  cond(J) max %.2f" % (worst, kap), flush=True)
