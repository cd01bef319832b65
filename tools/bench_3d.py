#!/usr/bin/env python3
"""3-D mapped mass matrices (trilinear hexahedra): Kronecker-sum kernel against the generic kernel, device resident, CUDA
events, best of 5.   python tools/bench_3d.py"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import interplib_b200 as ib  # noqa: E402
from interplib_b200 import DEVICE  # noqa: E402

GLL, GAUSS, LU = "lagrange-gauss-lobatto", "gauss", "lagrange-uniform"


def main():
    peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
    ctx = ib.Context(0)
    rng = np.random.default_rng(1)
    corners = np.array([[[2 * i0 - 1, 2 * i1 - 1, 2 * i2 - 1] for i0 in range(2) for i1 in range(2) for i2 in range(2)]], dtype=float)
    for p in (1, 2, 3, 4, 5):
        basis, rules, gb = [(GLL, p)] * 3, [(GAUSS, p + 1)] * 3, [(LU, 1)] * 3
        N = (p + 1) ** 3
        for kron in (True, False):
            E = max(256, int((8e9 if kron else 2e8) / (8 * N * N)) // 256 * 256)
            g = np.ascontiguousarray((corners + rng.uniform(-0.2, 0.2, (E, 8, 3))).transpose(0, 2, 1))
            d_g, d_M = ctx.alloc(g.nbytes).upload(g), ctx.alloc(E * N * N * 8)
            ctx.set_kron(kron)
            best = 1e30
            for i in range(6):
                ctx.timer_start(2)
                ctx.mass_matrix_mapped(basis, basis, rules, gb, d_g.ptr, 3, out=d_M.ptr, where=DEVICE, n_elements=E)
                ctx.timer_stop(2)
                ms = ctx.timer_ms(2)
                if i >= 1:
                    best = min(best, ms)
            ctx.set_kron(True)
            gbs = 8 * (24 + N * N) * E / (best * 1e-3) / 1e9
            print(json.dumps({"p": p, "N": N, "kernel": "mass3d_kron" if kron else "generic (geometry + mass)", "elements": E,
                              "ms": round(best, 4), "elements_per_s": E / (best * 1e-3), "achieved_gbs": round(gbs, 1),
                              "frac_of_hbm_peak": round(gbs / peak, 3)}), flush=True)
            d_g.free()
            d_M.free()
    ctx.close()


if __name__ == "__main__":
    main()
