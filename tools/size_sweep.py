#!/usr/bin/env python3
"""Device-resident mass-matrix kernel time against launch size (C5 shape, with and without the low-order kernel): the
rate must not depend on the number of elements per launch.   python tools/size_sweep.py"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import interplib_b200 as ib  # noqa: E402
from interplib_b200 import DEVICE, synthetic as syn  # noqa: E402

ctx = ib.Context(0)
cfg = syn.CONFIGS["c5"]
basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
Emax = 1 << 22
geom = syn.geometry(cfg, first=0, count=Emax)
d_g = ctx.alloc(geom.nbytes).upload(geom)
d_M = ctx.alloc(Emax * 625 * 8)
for lo in (True, False):
    ctx.set_lo_kernel(lo)
    for E in (1 << 18, 1 << 20, 3 << 19, 1 << 21, 3 << 20, 1 << 22):
        best = 1e9
        for rep in range(5):
            ctx.timer_start(3)
            ctx.mass_matrix_mapped(basis, basis, rules, gb, d_g.ptr, 2, out=d_M.ptr, where=DEVICE, n_elements=E)
            ctx.timer_stop(3)
            best = min(best, ctx.timer_ms(3))
        print(f"lo={lo} E={E:8d} {best:8.3f} ms  {E * 5064 / best / 1e6:8.1f} GB/s")
