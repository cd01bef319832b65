#!/usr/bin/env python3
"""Write rate of the mapped mass-matrix call (bilinear quads -> Kronecker kernel) over the order p: device resident, CUDA
events, best of 5.  Odd basis counts make every 256-byte warp store start off a 32-byte sector.   python tools/kron_sweep.py"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import interplib_b200 as ib  # noqa: E402
from interplib_b200 import DEVICE, synthetic as syn  # noqa: E402

GLL, GAUSS, LU = "lagrange-gauss-lobatto", "gauss", "lagrange-uniform"


def main():
    orders = [int(a) for a in sys.argv[1:]] or [7, 8, 11, 12, 14, 15, 16, 17, 19, 20]
    peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
    ctx = ib.Context(0)
    for p in orders:
        basis, rules, gb = [(GLL, p)] * 2, [(GAUSS, p + 1)] * 2, [(LU, 1)] * 2
        N = (p + 1) ** 2
        E = max(1024, int(24e9 / (8 * N * N)) // 1024 * 1024)
        side = int(np.ceil(np.sqrt(E)))
        geom = np.ascontiguousarray(syn.quad_mesh_geometry(side * side, seed=1).reshape(-1, 2, 4)[:E])
        d_g, d_M = ctx.alloc(geom.nbytes).upload(geom), ctx.alloc(E * N * N * 8)
        best = 1e30
        for i in range(7):
            ctx.timer_start(2)
            ctx.mass_matrix_mapped(basis, basis, rules, gb, d_g.ptr, 2, out=d_M.ptr, where=DEVICE, n_elements=E)
            ctx.timer_stop(2)
            ms = ctx.timer_ms(2)
            if i >= 2:
                best = min(best, ms)
        gbs = 8 * (8 + N * N) * E / (best * 1e-3) / 1e9
        print(json.dumps({"p": p, "N": N, "row_bytes": 8 * N, "rows_sector_aligned": (8 * N) % 32 == 0, "elements": E, "ms": round(best, 4),
                          "elements_per_s": E / (best * 1e-3), "achieved_gbs": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 3)}), flush=True)
        d_g.free()
        d_M.free()
    ctx.close()


if __name__ == "__main__":
    main()
