#!/usr/bin/env python3
"""Device-resident timings of the primal L2 projection (ib200_l2_project_mapped_batch) by fast diagonalisation against the
assemble-and-factorise path, BASELINE shapes C5 / C2-space / C3-order / C4: CUDA events on the context stream, 3 warm-ups,
best of 5; one JSON line per case.   python tools/bench_fd.py [--only c5]"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import interplib_b200 as ib  # noqa: E402
from interplib_b200 import DEVICE, synthetic as syn  # noqa: E402

GLL, GAUSS, LU = "lagrange-gauss-lobatto", "gauss", "lagrange-uniform"
CASES = {  # name: (p, rule order, elements for the fast path, elements for the factorisation path)
    "c5": (4, 5, 1 << 22, 1 << 20),
    "p6": (6, 7, 1 << 20, 1 << 17),
    "p8": (8, 9, 1 << 19, 1 << 14),
    "c4": (16, 17, 1 << 16, 1 << 11),
}


def peak():
    p = ROOT / "MEASURED_PEAKS.json"
    return float(json.loads(p.read_text())["hbm_gbs"]) if p.exists() else 6650.0


def timed(ctx, fn, reps=5, warm=3):
    best = 1e30
    for i in range(warm + reps):
        ctx.timer_start(2)
        fn()
        ctx.timer_stop(2)
        ms = ctx.timer_ms(2)
        if i >= warm:
            best = min(best, ms)
    return best


def main():
    only = sys.argv[sys.argv.index("--only") + 1] if "--only" in sys.argv else None
    fd_only = "--fd-only" in sys.argv
    ctx = ib.Context(0)
    rng = np.random.default_rng(0)
    for name, (p, ro, E_fd, E_ch) in CASES.items():
        if only and name != only:
            continue
        basis, rules, gb = [(GLL, p)] * 2, [(GAUSS, ro)] * 2, [(LU, 1)] * 2
        N, Q = (p + 1) ** 2, (ro + 1) ** 2
        for fd, E in ((True, E_fd), (False, E_ch)):
            if fd_only and not fd:
                continue
            side = int(np.ceil(np.sqrt(E)))
            geom = syn.quad_mesh_geometry(side * side, seed=1).reshape(-1, 2, 4)[:E]
            base = rng.standard_normal((min(E, 1 << 16), Q))
            f = np.ascontiguousarray(np.tile(base, (E // len(base) + 1, 1))[:E])
            d_g, d_f, d_u = ctx.alloc(geom.nbytes).upload(np.ascontiguousarray(geom)), ctx.alloc(f.nbytes).upload(f), ctx.alloc(E * N * 8)
            ctx.set_fastdiag(fd)
            ms = timed(ctx, lambda: ctx.l2_project_mapped(basis, rules, gb, d_g.ptr, 2, d_f.ptr, out=d_u.ptr, where=DEVICE, n_elements=E))
            ctx.set_fastdiag(True)
            bpe = 8 * (8 + Q + N)
            gbs = bpe * E / (ms * 1e-3) / 1e9
            print(json.dumps({"case": name, "solver": "fast diagonalisation" if fd else "assembly + Cholesky in scratch", "p": p,
                              "elements": E, "ms": round(ms, 4), "elements_per_s": E / (ms * 1e-3),
                              "algorithmic_bytes_per_element": bpe, "achieved_gbs": round(gbs, 1),
                              "frac_of_hbm_peak": round(gbs / peak(), 3)}), flush=True)
            for b in (d_g, d_f, d_u):
                b.free()
    ctx.close()


if __name__ == "__main__":
    main()
