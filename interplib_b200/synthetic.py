"""Synthetic inputs of the BASELINE.json configurations.

All generators are seeded ``numpy.random.default_rng`` streams so that tests, goldens and the bench see the
same data.  Layouts are the C-ABI ones (element index slowest).
"""
from __future__ import annotations

import dataclasses
import math

import numpy as np


@dataclasses.dataclass(frozen=True)
class Config:
    """One BASELINE.json configuration."""

    name: str
    ndim: int
    n_elements: int
    basis: tuple  # ((type, order), ...) of the function space
    rules: tuple  # ((method, order), ...)
    geom_basis: tuple
    n_phys: int
    description: str

    @property
    def N(self) -> int:
        return int(np.prod([o + 1 for _, o in self.basis]))

    @property
    def Q(self) -> int:
        return int(np.prod([o + 1 for _, o in self.rules]))


GLL = "lagrange-gauss-lobatto"


def _cfg2d(name, E, p, m, desc, basis=None):
    b = basis or ((GLL, p), (GLL, p))
    return Config(name, 2, E, tuple(b), (("gauss", m - 1), ("gauss", m - 1)), (("lagrange-uniform", 1),) * 2, 2, desc)


CONFIGS = {
    "c1": Config("c1", 1, 1024, ((GLL, 4),), (("gauss", 7),), (("bernstein", 1),), 1,
                 "1D mesh1d: 1,024 elements, Lagrange p=4 on GLL nodes, 8-pt Gauss rule; mass matrix + L2 projection of sin(x)"),
    "c2": _cfg2d("c2", 1 << 20, 6, 8, "2D quads: 1M elements, p=6, basis + gradient eval at 8x8 Gauss points, FP64"),
    "c3": _cfg2d("c3", 1 << 18, 8, 10, "2D 1-form basis mass matrices: 256K elements, p=8 (components (7,8) and (8,7)), FP64",
                 basis=((GLL, 7), (GLL, 8))),
    "c4": _cfg2d("c4", 1 << 16, 16, 18, "High-order 2D L2 projection: 64K elements, p=16, mass + L2 RHS"),
    "c5": _cfg2d("c5", 1 << 24, 4, 6, "Strong-scaling sweep: 16M quads, p=4, mass matrix + L2 projection"),
}

# the two component spaces of the 2-D 1-form basis of config 3 (de_rham_sequence.h:93-95: (p-1,p) and (p,p-1))
C3_SPACES = (((GLL, 7), (GLL, 8)), ((GLL, 8), (GLL, 7)))


def quad_mesh_geometry(n_elements: int, seed: int = 1, first: int = 0, count: int | None = None) -> np.ndarray:
    """Bilinear quads of a jittered structured grid on [0,1]^2 -> geom_dofs[count][2][2][2].

    ``geom[e][c][i0][i1]`` is coordinate ``c`` of the corner at reference position (xi0, xi1) =
    (2*i0-1, 2*i1-1): the ``lagrange-uniform`` order-1 coordinate DoFs that ``Quad.from_corners`` produces
    (reference python/interplib/domains.py:236-331).  Interior vertices are displaced by U(-0.2h, 0.2h).
    ``first``/``count`` select a contiguous slab of elements without generating the others' DoFs (the vertex
    grid is always generated whole so slabs are consistent)."""
    side = int(round(math.sqrt(n_elements)))
    if side * side != n_elements:
        raise ValueError("n_elements must be a perfect square")
    count = n_elements - first if count is None else count
    h = 1.0 / side
    rng = np.random.default_rng(seed)
    xs = np.linspace(0.0, 1.0, side + 1)
    vx, vy = np.meshgrid(xs, xs, indexing="ij")
    jit = rng.uniform(-0.2 * h, 0.2 * h, size=(2, side + 1, side + 1))
    jit[:, 0, :] = 0
    jit[:, -1, :] = 0
    jit[:, :, 0] = 0
    jit[:, :, -1] = 0
    vx = vx + jit[0]
    vy = vy + jit[1]
    e = np.arange(first, first + count)
    ex, ey = e // side, e % side
    geom = np.empty((count, 2, 2, 2))
    for i0 in range(2):
        for i1 in range(2):
            geom[:, 0, i0, i1] = vx[ex + i0, ey + i1]
            geom[:, 1, i0, i1] = vy[ex + i0, ey + i1]
    return geom


def line_mesh_geometry(n_elements: int, seed: int = 1) -> np.ndarray:
    """Segments of [0, 2*pi] with jittered break points -> geom_dofs[E][1][2] (Bernstein order-1 DoFs =
    the two end points, reference python/interplib/domains.py:200-233)."""
    rng = np.random.default_rng(seed)
    h = 2 * math.pi / n_elements
    knots = np.linspace(0.0, 2 * math.pi, n_elements + 1)
    knots[1:-1] += rng.uniform(-0.2 * h, 0.2 * h, n_elements - 1)
    geom = np.empty((n_elements, 1, 2))
    geom[:, 0, 0] = knots[:-1]
    geom[:, 0, 1] = knots[1:]
    return geom


def geometry(cfg: Config, first: int = 0, count: int | None = None, seed: int = 1) -> np.ndarray:
    """geom_dofs[count][n_phys][Ng] for a slab of the configuration's mesh."""
    if cfg.ndim == 1:
        g = line_mesh_geometry(cfg.n_elements, seed)
        count = cfg.n_elements - first if count is None else count
        return np.ascontiguousarray(g[first : first + count])
    return quad_mesh_geometry(cfg.n_elements, seed, first, count).reshape(-1, 2, 4)


def dofs(cfg: Config, count: int, seed: int = 2) -> np.ndarray:
    """U(0,1) degrees of freedom [count][N]."""
    return np.random.default_rng(seed).random((count, cfg.N))


def target_function(ndim: int):
    """sin(x) in 1-D, sin(x) cos(y) in 2-D."""
    if ndim == 1:
        return lambda x: np.sin(x)
    return lambda x, y: np.sin(x) * np.cos(y)
