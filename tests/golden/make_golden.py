#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the UNMODIFIED compiled reference (oracle/_ref, built from
/root/reference by oracle/build_ref.py) on sub-samples of the BASELINE.json configurations.

Run in the build container only (the reference sources do not exist on the GPU box):
    python tests/golden/make_golden.py
Each file holds the inputs (seeded synthetic data of interplib_b200/synthetic.py) and the reference's outputs in
the C-ABI batched layouts.  The reference publishes no golden vectors of its own, so
these are "outputs of the reference itself run here".
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from interplib_b200 import synthetic as syn  # noqa: E402
from oracle import refdrive as rd  # noqa: E402

OUT = Path(__file__).resolve().parent


def sample_elements(cfg, count):
    """`count` elements spread over the mesh (first, last and evenly spaced in between)."""
    idx = np.unique(np.linspace(0, cfg.n_elements - 1, count).astype(np.int64))
    geom = np.concatenate([syn.geometry(cfg, int(i), 1) for i in idx])
    return idx, geom


def mapped_fvals(cfg, coords):
    f = syn.target_function(cfg.ndim)
    return f(*[coords[:, a, 0, :] for a in range(cfg.n_phys)])


def make(cfg, count, with_mass=True, mass_stride=1, extra=None, full_mass_of=None):
    idx, geom = sample_elements(cfg, count)
    E = len(idx)
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    coords, det, jinv = rd.geometry(gb, rules, geom, cfg.n_phys)
    dofs = syn.dofs(cfg, E)
    masks = [0] + [1 << k for k in range(cfg.ndim)]
    data = dict(elements=idx, geom=geom, coords=coords, det=det, jinv=jinv, dofs=dofs,
                recon_masks=np.array(masks), recon=rd.reconstruct(basis, rules, dofs, masks))
    fvals = mapped_fvals(cfg, coords)
    data["fvals"] = fvals
    data["l2_dual"] = rd.l2_dual(basis, rules, fvals, gb, geom, cfg.n_phys)
    if with_mass:
        m = rd.mass(basis, basis, rules, E, gb, geom, cfg.n_phys)
        data["mass_stride"] = np.array(mass_stride)
        data["mass"] = m.reshape(E, -1)[:, ::mass_stride].copy()
        if full_mass_of is not None:
            data["mass_full"] = m[full_mass_of].copy()  # one whole matrix next to the strided samples
            data["mass_full_of"] = np.array(full_mass_of)
        data["mass_ref_domain"] = rd.mass(basis, basis, rules, 1).reshape(1, -1)[:, ::mass_stride].copy()
        data["primal"] = np.stack([np.linalg.solve(m[e], data["l2_dual"][e]) for e in range(E)])
    if extra:
        extra(data, geom, E)
    np.savez(OUT / f"{cfg.name}.npz", **data)
    print(cfg.name, {k: v.shape for k, v in data.items()})


def c3_blocks(data, geom, E):
    cfg = syn.CONFIGS["c3"]
    rules, gb = list(cfg.rules), list(cfg.geom_basis)
    for a, sa in enumerate(syn.C3_SPACES):
        for b, sb in enumerate(syn.C3_SPACES):
            # block M_ab = compute_mass_matrix(space_b, space_a, smap): rows = space a (out), cols = space b (in)
            data[f"block_{a}{b}"] = rd.mass(list(sb), list(sa), rules, E, gb, geom, 2)


def c5_gradient(data, geom, E):
    cfg = syn.CONFIGS["c5"]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    lower = [(t, o - 1) for t, o in basis]
    for i in range(2):
        for j in range(2):
            data[f"gradmass_{i}{j}"] = rd.mass(basis, basis, rules, E, gb, geom, 2, idx_in=i, idx_out=j)
            data[f"gradmass_mixed_{i}{j}"] = rd.mass(basis, lower, rules, E, gb, geom, 2, idx_in=i, idx_out=j)


if __name__ == "__main__":
    only = sys.argv[1:]
    if only == ["c4"]:
        make(syn.CONFIGS["c4"], 2, mass_stride=61, full_mass_of=1)
        sys.exit(0)
    make(syn.CONFIGS["c1"], 16)
    make(syn.CONFIGS["c2"], 8)
    make(syn.CONFIGS["c3"], 3, with_mass=False, extra=c3_blocks)
    make(syn.CONFIGS["c4"], 2, mass_stride=61, full_mass_of=1)
    make(syn.CONFIGS["c5"], 8, extra=c5_gradient)
