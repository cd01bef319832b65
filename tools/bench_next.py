#!/usr/bin/env python3
"""Device-resident timings of the SURVEY 8f "next" rows (N1 primal solve, N2 arbitrary points, N3 derivative DoFs) on one
GPU: CUDA events on the context stream, 3 warm-ups, best of 5; prints one JSON line per kernel with the algorithmic
bytes per element and the achieved fraction of the measured HBM copy bandwidth.   python tools/bench_next.py"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import interplib_b200 as ib  # noqa: E402
from interplib_b200 import DEVICE, synthetic as syn  # noqa: E402


def peak():
    p = ROOT / "MEASURED_PEAKS.json"
    return float(json.loads(p.read_text())["hbm_gbs"]) if p.exists() else 6650.0


def timed(ctx, fn, reps=5, warm=3, setup=None):
    best = 1e30
    for i in range(warm + reps):
        if setup:
            setup()
        ctx.timer_start(2)
        fn()
        ctx.timer_stop(2)
        ms = ctx.timer_ms(2)
        if i >= warm:
            best = min(best, ms)
    return best


def report(name, E, bytes_per_el, ms):
    gbs = bytes_per_el * E / (ms * 1e-3) / 1e9
    print(json.dumps({"kernel": name, "elements": E, "ms": round(ms, 4), "elements_per_s": E / (ms * 1e-3),
                      "algorithmic_bytes_per_element": bytes_per_el, "achieved_gbs": round(gbs, 1),
                      "frac_of_hbm_peak": round(gbs / peak(), 3)}))


def main():
    ctx = ib.Context(0)
    rng = np.random.default_rng(0)
    # ---- N1: M u = b for the C5 system (25 x 25), 2^19 elements
    cfg = syn.CONFIGS["c5"]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    E = 1 << 19
    geom = syn.geometry(cfg, first=0, count=E)
    f = rng.standard_normal((E, cfg.Q))
    d_g, d_f = ctx.alloc(geom.nbytes).upload(geom), ctx.alloc(f.nbytes).upload(f)
    d_M, d_b = ctx.alloc(E * cfg.N * cfg.N * 8), ctx.alloc(E * cfg.N * 8)

    def assemble():
        ctx.l2_system_mapped(basis, rules, gb, d_g.ptr, 2, d_f.ptr, mass=d_M.ptr, rhs=d_b.ptr, where=DEVICE, n_elements=E)

    ms = timed(ctx, lambda: ctx.spd_solve(d_M.ptr, d_b.ptr, where=DEVICE, n=cfg.N, n_elements=E), setup=assemble)
    report("N1 spd_solve 25x25 (read M + b, write u)", E, 8 * (cfg.N * cfg.N + 2 * cfg.N), ms)
    for b in (d_g, d_f, d_M, d_b):
        b.free()
    # ---- N1 fused: the whole projection_l2_primal of C5 (geometry + f -> u; M assembled and solved in device scratch)
    E = 1 << 20
    geom = syn.geometry(cfg, first=0, count=E)
    f = rng.standard_normal((E, cfg.Q))
    d_g, d_f, d_u = ctx.alloc(geom.nbytes).upload(geom), ctx.alloc(f.nbytes).upload(f), ctx.alloc(E * cfg.N * 8)
    ms = timed(ctx, lambda: ctx.l2_project_mapped(basis, rules, gb, d_g.ptr, 2, d_f.ptr, out=d_u.ptr, where=DEVICE, n_elements=E))
    report("N1 l2_project C5 (geometry + f -> u), device resident", E, 8 * (geom.shape[1] * geom.shape[2] + cfg.Q + cfg.N), ms)
    for buf in (d_g, d_f, d_u):
        buf.free()
    import time

    u = np.empty((E, cfg.N))
    best = 1e30
    for _ in range(4):
        t0 = time.perf_counter()
        ctx.l2_project_mapped(basis, rules, gb, geom, 2, f, out=u)
        best = min(best, time.perf_counter() - t0)
    print(json.dumps({"kernel": "N1 l2_project C5 end to end (host arrays through the C-ABI)", "elements": E, "ms": round(best * 1e3, 3),
                      "elements_per_s": E / best, "h2d_bytes": geom.nbytes + f.nbytes, "d2h_bytes": u.nbytes}))
    best = 1e30
    for _ in range(2):
        t0 = time.perf_counter()
        M, b = ctx.l2_system_mapped(basis, rules, gb, geom, 2, f)
        ctx.spd_solve(M, b)
        best = min(best, time.perf_counter() - t0)
    print(json.dumps({"kernel": "N1 same projection as l2_system + spd_solve with M through the host (r01e flow)", "elements": E,
                      "ms": round(best * 1e3, 3), "elements_per_s": E / best}))
    del M, b, u, geom, f
    # ---- N1 at the C2 (49 x 49, shared-memory kernel) and C4 (289 x 289, in-place kernel) system sizes: random SPD matrices
    for n, E in ((49, 1 << 16), (289, 1 << 10)):
        G = rng.standard_normal((E, n, n))
        M = G @ np.swapaxes(G, 1, 2) + n * np.eye(n)
        b = rng.standard_normal((E, n))
        d_M, d_b = ctx.alloc(M.nbytes).upload(M), ctx.alloc(b.nbytes).upload(b)
        del G
        # the in-place kernel (n > 127) overwrites M with its factor: upload it again before every timed call
        restore = (lambda d_M=d_M, M=M: d_M.upload(M)) if n > 127 else None
        ms = timed(ctx, lambda: ctx.spd_solve(d_M.ptr, d_b.ptr, where=DEVICE, n=n, n_elements=E), setup=restore)
        report(f"N1 spd_solve {n}x{n} (read M + b, write u)", E, 8 * (n * n + 2 * n), ms)
        for buf in (d_M, d_b):
            buf.free()
    # ---- N2: reconstruct value + gradient at 64 arbitrary points per element, C2 space, 2^18 elements
    cfg = syn.CONFIGS["c2"]
    basis = list(cfg.basis)
    E, P = 1 << 18, 64
    u, x = rng.random((E, cfg.N)), rng.uniform(-1, 1, (E, 2, P))
    d_u, d_x, d_o = ctx.alloc(u.nbytes).upload(u), ctx.alloc(x.nbytes).upload(x), ctx.alloc(E * 3 * P * 8)
    ms = timed(ctx, lambda: ctx.reconstruct_points(basis, d_u.ptr, d_x.ptr, [0, 1, 2], out=d_o.ptr, where=DEVICE, n_elements=E,
                                                   n_points=P))
    report("N2 reconstruct_points p=6, 64 points, value+gradient", E, 8 * (cfg.N + 2 * P + 3 * P), ms)
    # ---- N2: FunctionSpace.evaluate -- all N basis values at 64 points per element, 2^16 elements (1.6 GB out)
    Ee = 1 << 16
    d_phi = ctx.alloc(Ee * P * cfg.N * 8)
    ms = timed(ctx, lambda: ctx.evaluate_points(basis, d_x.ptr, out=d_phi.ptr, where=DEVICE, n_elements=Ee, n_points=P))
    report("N2 evaluate_points p=6, 64 points (basis values [E][P][N])", Ee, 8 * (2 * P + P * cfg.N), ms)
    d_phi.free()
    # ---- N3: derivative DoFs along dimension 0 and 1, C2 space, 2^20 elements
    for b in (d_u, d_x, d_o):
        b.free()
    E = 1 << 20
    u = rng.random((E, cfg.N))
    d_u, d_o = ctx.alloc(u.nbytes).upload(u), ctx.alloc(E * 42 * 8)
    for k in range(2):
        ms = timed(ctx, lambda: ctx.dof_derivative(basis, d_u.ptr, k, out=d_o.ptr, where=DEVICE, n_elements=E))
        report(f"N3 dof_derivative p=6 along dim {k}", E, 8 * (49 + 42), ms)
    ctx.close()


if __name__ == "__main__":
    main()
