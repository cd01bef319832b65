"""The reference's CPU path timed on this host's cores.  TEST / BENCH INFRASTRUCTURE (used only by bench.py's
``--impl reference`` arm and its ``cpu_baseline`` leg, after the GPU timed region).

``reference`` = the UNMODIFIED compiled interplib (oracle/_ref) driven per element through its own Python API
(oracle/refdrive.py).  The reference holds the GIL and has no parallel region, so "all host cores" means one worker
process per core.  Workers are *spawned* (fresh interpreters that exit normally), not forked, and never import the
repo's own shared library: a process-exit hook then sees exactly the reference's ``_interp*.so`` mapped.
``port`` = the C restatement (oracle/interp_oracle.c, OpenMP) when the compiled reference is absent.

Workloads follow interplib_b200/synthetic.py; importing that module does not load libinterp_b200.so.
"""
from __future__ import annotations

import os
import time

import numpy as np

from interplib_b200 import synthetic as syn


def slab(cfg, first: int, count: int) -> np.ndarray:
    """geom_dofs of `count` elements starting at `first`, wrapping around the configuration's mesh."""
    first %= cfg.n_elements
    if cfg.ndim == 2 and first + count <= cfg.n_elements:
        return syn.geometry(cfg, first=first, count=count)
    if cfg.ndim == 1:
        return np.ascontiguousarray(syn.geometry(cfg)[(first + np.arange(count)) % cfg.n_elements])
    parts, left = [], count
    while left > 0:
        n = min(left, cfg.n_elements - first)
        parts.append(syn.geometry(cfg, first=first, count=n))
        first, left = 0, left - n
    return np.concatenate(parts)


def ref_worker(args) -> float:
    """Run the compiled reference per element on a slab; returns the seconds spent inside the reference calls."""
    name, first, count, seed = args
    from oracle import refdrive as rd

    cfg = syn.CONFIGS[name]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    rng = np.random.default_rng(seed)
    if name == "c2":
        dofs = rng.random((count, cfg.N))
        t0 = time.perf_counter()
        rd.reconstruct(basis, rules, dofs, [0, 1, 2])
        return time.perf_counter() - t0
    geom = slab(cfg, first, count)
    fvals = rng.standard_normal((count, cfg.Q))
    t0 = time.perf_counter()
    if name == "c3":
        for sa in syn.C3_SPACES:
            for sb in syn.C3_SPACES:
                rd.mass(list(sb), list(sa), rules, count, gb, geom, cfg.n_phys)
    else:
        rd.mass(basis, basis, rules, count, gb, geom, cfg.n_phys)
        rd.l2_dual(basis, rules, fvals, gb, geom, cfg.n_phys)
    return time.perf_counter() - t0


def port_step(name: str, count: int, seed: int) -> float:
    from oracle import oracle as orc

    cfg = syn.CONFIGS[name]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    rng = np.random.default_rng(seed)
    if name == "c2":
        dofs = rng.random((count, cfg.N))
        t0 = time.perf_counter()
        orc.reconstruct(basis, rules, dofs, [0, 1, 2])
        return time.perf_counter() - t0
    geom = slab(cfg, 0, count)
    f = rng.standard_normal((count, cfg.Q))
    t0 = time.perf_counter()
    _, det, _ = orc.geometry(gb, rules, geom, cfg.n_phys, want_coords=False, want_jinv=False)
    if name == "c3":
        for sa in syn.C3_SPACES:
            for sb in syn.C3_SPACES:
                orc.mass(list(sb), list(sa), rules, count, det=det)
    else:
        orc.mass(basis, basis, rules, count, det=det)
        orc.l2_dual(basis, rules, f, det=det)
    return time.perf_counter() - t0


def _init_worker():
    # one BLAS/OpenMP thread per worker: the workers are the parallelism
    os.environ["OMP_NUM_THREADS"] = "1"
    from oracle import refdrive as rd

    rd.ref()  # import the reference now, outside the timed calls


class CpuArm:
    """`reference` (compiled interplib, one spawned process per core) or `port` (C oracle, OpenMP)."""

    def __init__(self, name: str, single_thread: bool = False):
        from oracle import build_ref

        self.name = name
        self.kind = "reference" if build_ref.built() else "port"
        self.cores = 1 if single_thread else (os.cpu_count() or 1)
        self.pool = None
        if self.kind == "reference":
            if self.cores > 1:
                import multiprocessing as mp

                self.pool = mp.get_context("spawn").Pool(self.cores, initializer=_init_worker)
                self.pool.map(_warm, range(self.cores))
        else:
            from oracle import oracle as orc

            orc.build()
            if single_thread:
                os.environ["OMP_NUM_THREADS"] = "1"
            self.cores = 1 if single_thread else orc.threads()

    def run(self, count: int, seed: int = 0) -> float:
        """Process `count` elements; returns the seconds spent in the reference's (or the port's) own calls -- the slowest
        worker's when several run side by side.  Generating the synthetic inputs (the jittered vertex grid of a 16M-quad mesh
        costs more than a small sample's arithmetic) is not charged to the reference."""
        if self.kind == "port":
            return port_step(self.name, count, seed)
        if self.pool is None:
            return ref_worker((self.name, 0, count, seed))
        per = max(1, count // self.cores)
        jobs = [(self.name, i * per, per, seed + i) for i in range(self.cores)]
        return max(self.pool.map(ref_worker, jobs, chunksize=1))

    def effective_count(self, count: int) -> int:
        if self.pool is None:
            return count
        return max(1, count // self.cores) * self.cores

    def calibrate(self, target_s: float, cap: int = 1 << 22) -> int:
        """Elements per step so that one step takes about target_s (at most `cap`)."""
        n = self.effective_count(max(self.cores * 4, 8))
        self.run(n)  # first call: imports, table caches, page faults -- not representative
        t = self.run(n)
        while t < 0.5 and n < cap:
            n *= 4
            t = self.run(n)
        return self.effective_count(max(self.cores, min(cap, int(n * target_s / max(t, 1e-9)))))

    def close(self):
        if self.pool:
            self.pool.close()  # workers leave their loop and exit normally (atexit hooks run)
            self.pool.join()
            self.pool = None


def _warm(_):
    return os.getpid()
