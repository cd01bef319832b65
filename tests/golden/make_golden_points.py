#!/usr/bin/env python3
"""Generate tests/golden/points.npz and incidence.npz -- arbitrary-point evaluation and
incidence matrices / derivative DoFs (row N3: incidence_matrix, DegreesOfFreedom.derivative) by the UNMODIFIED compiled
reference (oracle/_ref): FunctionSpace.evaluate, interplib.reconstruct and BasisSpecs.values/.derivatives on seeded
random points.  Run in the build container only:  python tests/golden/make_golden_points.py"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import refdrive as rd  # noqa: E402

OUT = Path(__file__).resolve().parent
ALL_BASIS = ["legendre", "lagrange-gauss-lobatto", "lagrange-gauss", "lagrange-uniform", "lagrange-chebyshev-gauss", "bernstein"]
# name -> tensor-product space; c2/c4 are the BASELINE configurations' spaces
SPACES = {
    "c2": [("lagrange-gauss-lobatto", 6)] * 2,
    "c4": [("lagrange-gauss-lobatto", 16)] * 2,
    "line": [("bernstein", 3)],
    "mixed2": [("legendre", 3), ("lagrange-chebyshev-gauss", 2)],
    "mixed3": [("lagrange-uniform", 2), ("bernstein", 1), ("lagrange-gauss", 3)],
}


DERIV_SPACES = {
    "c2": [("lagrange-gauss-lobatto", 6)] * 2,
    "c3a": [("lagrange-gauss-lobatto", 7), ("lagrange-gauss-lobatto", 8)],
    "line": [("bernstein", 3)],
    "mixed3": [("legendre", 4), ("bernstein", 2), ("lagrange-chebyshev-gauss", 3)],
}


def main():
    rng = np.random.default_rng(2024)
    data = {}
    for name, basis in SPACES.items():
        E, P = 3, 7
        N = int(np.prod([o + 1 for _, o in basis]))
        x = rng.uniform(-1.0, 1.0, (E, len(basis), P))
        dofs = rng.random((E, N))
        data[f"{name}_x"] = x
        data[f"{name}_dofs"] = dofs
        data[f"{name}_eval"] = rd.evaluate_points(basis, x)
        data[f"{name}_recon"] = rd.reconstruct_points(basis, dofs, x)
    x1 = rng.uniform(-1.0, 1.0, 9)
    data["x1"] = x1
    for bt in ALL_BASIS:
        for order in (1, 4, 9):
            v, d = rd.basis_points_1d((bt, order), x1)
            data[f"val_{bt}_{order}"] = v
            data[f"der_{bt}_{order}"] = d
    np.savez(OUT / "points.npz", **data)
    print({k: v.shape for k, v in data.items()})

    # N3: incidence matrices and DegreesOfFreedom.derivative
    inc = {}
    for bt in ALL_BASIS:
        for order in (1, 2, 5, 8, 16):
            inc[f"inc_{bt}_{order}"] = rd.incidence_matrix((bt, order))
    for name, basis in DERIV_SPACES.items():
        N = int(np.prod([o + 1 for _, o in basis]))
        dofs = rng.standard_normal((4, N))
        inc[f"{name}_dofs"] = dofs
        for k in range(len(basis)):
            inc[f"{name}_d{k}"] = rd.dof_derivative(basis, dofs, k)
    np.savez(OUT / "incidence.npz", **inc)
    print({k: v.shape for k, v in inc.items() if not k.startswith("inc_")})


if __name__ == "__main__":
    main()
