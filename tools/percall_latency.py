"""Cost of ONE call through the per-element (reference-style) API: this repository's drop-in (`dropin/interplib`, Python +
ctypes over the C ABI, one kernel launch and two PCIe copies per call) against the compiled reference (`oracle/_ref`, a
CPython extension looping on the host), at the element shapes of BASELINE configs 1, 5 and 2.

The same script body runs in two fresh interpreters that differ only in which `interplib` is first on the path (both
cannot live in one process).  Output: one JSON object, by default into profiles/r03_percall.json.

    python tools/percall_latency.py [--out profiles/r03_percall.json]

Reference entry points timed (src/python/module.c:876-898 registers them): compute_mass_matrix
(src/python/mass_matrices.c:161-339), DegreesOfFreedom.reconstruct_at_integration_points
(src/python/degrees_of_freedom.c:426-510), SpaceMap.__new__ (src/python/mappings.c:304-556).
"""
import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]

INNER = r"""
import json, sys, time
import numpy as np
import interplib as ip

SHAPES = {
    "c1 (1-D, p=4, 8-pt Gauss)": dict(basis=[("lagrange-gauss-lobatto", 4)], rule=[7], gb=[("lagrange-uniform", 1)]),
    "c5 (2-D, p=4, 6x6 Gauss)": dict(basis=[("lagrange-gauss-lobatto", 4)] * 2, rule=[5, 5], gb=[("lagrange-uniform", 1)] * 2),
    "c2 (2-D, p=6, 8x8 Gauss)": dict(basis=[("lagrange-gauss-lobatto", 6)] * 2, rule=[7, 7], gb=[("lagrange-uniform", 1)] * 2),
}

def best_us(fn, seconds=0.4):
    fn(); fn()
    n, best = 1, float("inf")
    t_end = time.perf_counter() + seconds
    while time.perf_counter() < t_end:
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        dt = (time.perf_counter() - t0) / n
        best = min(best, dt)
        if dt * n < 2e-3:
            n *= 2
    return best * 1e6

out = {}
rng = np.random.default_rng(3)
for name, s in SHAPES.items():
    d = len(s["basis"])
    fs = ip.FunctionSpace(*(ip.BasisSpecs(t, o) for t, o in s["basis"]))
    gs = ip.FunctionSpace(*(ip.BasisSpecs(t, o) for t, o in s["gb"]))
    isp = ip.IntegrationSpace(*(ip.IntegrationSpecs(o) for o in s["rule"]))
    # a mildly distorted unit cell: corner DoFs of the order-1 coordinate maps
    corners = [np.array(np.meshgrid(*([[0.0, 1.0]] * d), indexing="ij")[k]) + 0.05 * rng.random([2] * d) for k in range(d)]
    gdofs = [ip.DegreesOfFreedom(gs, c) for c in corners]
    maps = [ip.CoordinateMap(g, isp) for g in gdofs]
    smap = ip._interp.SpaceMap(*maps)
    dof = ip.DegreesOfFreedom(fs, rng.random([o + 1 for _, o in s["basis"]]))
    out[name] = {
        "compute_mass_matrix(fs, fs, smap)": best_us(lambda: ip.compute_mass_matrix(fs, fs, smap)),
        "dof.reconstruct_at_integration_points(isp)": best_us(lambda: dof.reconstruct_at_integration_points(isp)),
        "CoordinateMap x d + SpaceMap": best_us(lambda: ip._interp.SpaceMap(*[ip.CoordinateMap(g, isp) for g in gdofs])),
    }
print("PERCALL " + json.dumps(out))
"""


def run(first_on_path: Path) -> dict:
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([str(first_on_path), str(ROOT)])
    env["INTERP_B200_REPO"] = str(ROOT)
    p = subprocess.run([sys.executable, "-c", INNER], env=env, capture_output=True, text=True, timeout=600, cwd="/tmp")
    for line in p.stdout.splitlines():
        if line.startswith("PERCALL "):
            return json.loads(line[8:])
    raise RuntimeError(f"{first_on_path}: no result\n{p.stdout[-2000:]}\n{p.stderr[-2000:]}")


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=str(ROOT / "profiles" / "r03_percall.json"))
    args = ap.parse_args()
    ref = run(ROOT / "oracle" / "_ref")
    ours = run(ROOT / "dropin")
    table = {}
    for shape in ref:
        table[shape] = {call: {"reference_us": round(ref[shape][call], 2), "dropin_us": round(ours[shape][call], 2),
                               "dropin_over_reference": round(ours[shape][call] / ref[shape][call], 2)}
                        for call in ref[shape]}
    rec = {"what": "best-of microseconds per single call through the per-element API, host arrays in and out",
           "dropin": "dropin/interplib -> interplib_b200 (ctypes) -> libinterp_b200.so: one launch + H2D + D2H per call",
           "reference": "oracle/_ref/interplib (the unmodified reference, compiled here)", "table": table}
    Path(args.out).write_text(json.dumps(rec, indent=1) + "\n")
    print(json.dumps(rec, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())
