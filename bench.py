#!/usr/bin/env python3
"""bench.py -- throughput of the interplib hot path on B200 (contract: see README / DESIGN.md section 6).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c5|c4|c3|c1] [--impl ours|reference]

One "step" = one pass of the hot path over one batch of synthetic elements.  The default workload is
BASELINE.json config 2 (2-D quads, 2^20 elements, p=6, basis + gradient evaluation at 8x8 Gauss points, FP64);
metric = elements (basis+grad evals) per second, whole job.  Elements are independent: under torchrun every rank
owns a GPU and its own slab of elements (weak scaling, no data-path collective); timing is on the device (CUDA
events on the launching stream), max over ranks.  Rank 0 prints ONE JSON line.

`--impl reference` times the reference's own CPU implementation (oracle/_ref: the unmodified interplib extension
driven per element through its Python API, fanned out over all host cores; falls back to the C oracle port when
the compiled reference is absent) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from interplib_b200 import synthetic as syn  # noqa: E402

# ncu-measured DRAM traffic of the dominant kernel (dram__bytes_read.sum + dram__bytes_write.sum of one launch from the
# `ncu --set full` captures summarised under profiles/*_full.txt) as (bytes per launch, elements in that launch); the
# bench line scales it to the elements of its own launch.  Slightly below the algorithmic bytes where the tail of the
# output is still dirty in the 126 MB L2 when the kernel ends.
NCU_TRAFFIC_BYTES = {
    "c2": (1970064432, 1 << 20),   # profiles/r01e_c2_eval2d_full.txt
    "c5": (5249943208, 1 << 20),   # profiles/r01c_c5_mass2d_lo_full.txt
    "c4": (43748198408, 1 << 16),  # profiles/r01e_c4_mass2d_mma_full.txt
    "c3": (11733279000, 1 << 18),  # profiles/r01e_c3_mass2d_full.txt
    "c1": None,
}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------ workloads
class Workload:
    """Per-rank batch of one BASELINE configuration, device resident."""

    def __init__(self, name: str, ctx, rank: int, n_elements: int | None):
        self.name = name
        self.cfg = cfg = syn.CONFIGS[name]
        self.ctx = ctx
        self.basis, self.rules, self.gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
        E = n_elements or self.default_elements()
        self.E = E
        N, Q = cfg.N, cfg.Q
        rng = np.random.default_rng(2 + rank)
        self.blocks = None
        if name == "c2":
            # basis + gradient eval: dofs[E][49] -> out[E][3][64]
            self.h_in = {"dofs": rng.random((E, N))}
            self.out_shapes = {"out": (E, 3, Q)}
            self.bytes_per_elem = 8 * (N + 3 * Q)
            self.dominant = "eval2d_kernel (K1: value + gradient)"
            self.dominant_bytes_per_elem = self.bytes_per_elem
            self.flops_per_elem = 2 * (2 * 7 * 7 * 8 + 3 * 7 * 8 * 8)
            self.metric = "basis_grad_evals_per_s"
            self.unit = "elements/s"
        elif name == "c3":
            # the four determinant-weighted blocks M_ab = compute_mass_matrix(space_b, space_a, smap) of the 2-D 1-form
            # basis (SURVEY 8c caveat): geometry[E][2][4] -> 4 x M[E][72][72] (= the 144 x 144 element matrix, block-wise)
            first = (rank * E) % max(cfg.n_elements - E + 1, 1)
            self.h_in = {"geom": syn.geometry(cfg, first=first, count=E)}
            self.blocks = [(a, b, list(sa), list(sb)) for a, sa in enumerate(syn.C3_SPACES) for b, sb in enumerate(syn.C3_SPACES)]
            self.out_shapes = {f"block_{a}{b}": (E, 72, 72) for a, b, _, _ in self.blocks}
            self.bytes_per_elem = 8 * (8 + 4 * 72 * 72)
            self.dominant = "mass2d_kernel (K2, one 72x72 block per launch, 4 launches per step)"
            self.dominant_bytes_per_elem = 8 * (8 + 72 * 72)
            self.metric = "element_mass_matrices_per_s"
            self.unit = "elements/s"
        else:
            # mass matrix (+ L2 right-hand side): geometry[E][2][4] (+ f[E][Q]) -> M[E][N][N] (+ b[E][N])
            first = (rank * E) % max(cfg.n_elements - E + 1, 1)
            if cfg.ndim == 2:
                geom = syn.geometry(cfg, first=first, count=E)
            else:
                geom = _slab(cfg, 0, E)
            self.h_in = {"geom": geom, "fvals": rng.standard_normal((E, Q))}
            self.out_shapes = {"mass": (E, N, N), "rhs": (E, N)}
            self.bytes_per_elem = 8 * (geom[0].size + Q + N * N + N)
            self.dominant = ({"c5": "mass2d_lo_kernel (K2-lo, geometry fused)", "c4": "mass2d_mma_kernel (K2-hi DMMA, geometry fused)"}
                             .get(name, "mass2d_kernel (K2, geometry fused)") if cfg.ndim == 2 else "geometry + mass kernels (1-D generic)")
            self.dominant_bytes_per_elem = 8 * (geom[0].size + N * N)
            self.metric = "element_mass_matrices_per_s"
            self.unit = "elements/s"
        self.d_in = {k: ctx.alloc(v.nbytes).upload(v) for k, v in self.h_in.items()}
        self.d_out = {k: ctx.alloc(int(np.prod(s)) * 8) for k, s in self.out_shapes.items()}

    def default_elements(self) -> int:
        # the whole configuration per launch (C3: 2^18 elements, four 10.9 GB blocks; C4: 2^16 elements, 43.8 GB); C5's
        # 16M quads x 5 kB = 84 GB are executed in slabs of 2^22 elements reusing one output slab
        return {"c1": 1024, "c2": 1 << 20, "c3": 1 << 18, "c4": 1 << 16, "c5": 1 << 22}[self.name]

    @property
    def E2(self) -> int:
        """Elements of one end-to-end step: the host-pointer path streams its arrays through PCIe in 16 MB chunks, so its
        rate does not depend on the batch; C3-C5 use a slab of the device-resident batch to bound pinned host memory
        (C3's four blocks of 2^18 elements would be 43 GB of pinned buffers)."""
        return min(self.E, {"c3": 1 << 15, "c4": 1 << 13, "c5": 1 << 20}.get(self.name, self.E))

    def step_device(self, dominant_only: bool = False) -> None:
        from interplib_b200 import DEVICE

        c = self.ctx
        if self.name == "c2":
            c.reconstruct(self.basis, self.rules, self.d_in["dofs"].ptr, [0, 1, 2], out=self.d_out["out"].ptr, where=DEVICE,
                          n_elements=self.E)
        elif self.name == "c3":
            for a, b, sa, sb in (self.blocks[:1] if dominant_only else self.blocks):
                c.mass_matrix_mapped(sb, sa, self.rules, self.gb, self.d_in["geom"].ptr, 2, out=self.d_out[f"block_{a}{b}"].ptr,
                                     where=DEVICE, n_elements=self.E)
        else:
            if dominant_only:
                c.mass_matrix_mapped(self.basis, self.basis, self.rules, self.gb, self.d_in["geom"].ptr, self.cfg.n_phys,
                                     out=self.d_out["mass"].ptr, where=DEVICE, n_elements=self.E)
                return
            # mass matrix + L2 right-hand side of the same elements: one fused kernel at low order (K2-lo + K3),
            # two launches otherwise (ib200_l2_system_mapped_batch)
            c.l2_system_mapped(self.basis, self.rules, self.gb, self.d_in["geom"].ptr, self.cfg.n_phys, self.d_in["fvals"].ptr,
                               mass=self.d_out["mass"].ptr, rhs=self.d_out["rhs"].ptr, where=DEVICE, n_elements=self.E)

    # end to end through the public host-pointer API (pinned host buffers, H2D + D2H inside the call)
    def e2e_prepare(self) -> None:
        c = self.ctx
        n = self.E2
        self.p_in = {k: c.pinned((n,) + v.shape[1:]) for k, v in self.h_in.items()}
        for k, v in self.h_in.items():
            self.p_in[k].array[...] = v[:n]
        self.p_out = {k: c.pinned((n,) + tuple(s[1:])) for k, s in self.out_shapes.items()}

    def step_e2e(self) -> float:
        c = self.ctx
        if self.name == "c2":
            c.reconstruct(self.basis, self.rules, self.p_in["dofs"].array, [0, 1, 2], out=self.p_out["out"].array)
            return float(self.p_out["out"].array[-1, 0, 0])
        if self.name == "c3":
            for a, b, sa, sb in self.blocks:
                c.mass_matrix_mapped(sb, sa, self.rules, self.gb, self.p_in["geom"].array, 2, out=self.p_out[f"block_{a}{b}"].array)
            return float(self.p_out["block_11"].array[-1, 0, 0])
        c.l2_system_mapped(self.basis, self.rules, self.gb, self.p_in["geom"].array, self.cfg.n_phys, self.p_in["fvals"].array,
                           mass=self.p_out["mass"].array, rhs=self.p_out["rhs"].array)
        return float(self.p_out["rhs"].array[-1, 0])

    @property
    def h2d_bytes(self) -> int:
        return int(sum(v.nbytes // self.E * self.E2 for v in self.h_in.values()))

    @property
    def d2h_bytes(self) -> int:
        return int(sum(int(np.prod(s[1:])) * 8 * self.E2 for s in self.out_shapes.values()))


def bind_to_gpu_numa_node(gpu_index: int) -> str:
    """Pin this rank's threads to the CPUs next to its GPU (NVML's ideal affinity) before any pinned host buffer is
    allocated: the end-to-end path is PCIe-bound, and on a two-socket host a rank whose staging memory sits on the far socket
    shares the inter-socket link with every other rank.  Best effort; returns a note for the JSON line."""
    if os.environ.get("IB200_NO_NUMA_BIND"):
        return "disabled"
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return f"{len(os.sched_getaffinity(0))} CPUs (NVML ideal affinity of GPU {gpu_index})"
    except Exception as exc:  # noqa: BLE001 -- NVML missing or not permitted: run unbound
        return f"unbound ({type(exc).__name__})"


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self):
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self, gpu_indices) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for line in Path(self.path).read_text().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9 or not f[0].isdigit() or int(f[0]) not in gpu_indices:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU arms
def _slab(cfg, first: int, count: int) -> np.ndarray:
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


def _ref_worker(args):
    """Run the compiled reference per element on a slab (one process per core)."""
    name, first, count, seed = args
    from oracle import refdrive as rd

    cfg = syn.CONFIGS[name]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    rng = np.random.default_rng(seed)
    t0 = time.perf_counter()
    if name == "c2":
        rd.reconstruct(basis, rules, rng.random((count, cfg.N)), [0, 1, 2])
    else:
        geom = _slab(cfg, first, count)
        t0 = time.perf_counter()
        if name == "c3":
            for sa in syn.C3_SPACES:
                for sb in syn.C3_SPACES:
                    rd.mass(list(sb), list(sa), rules, count, gb, geom, cfg.n_phys)
        else:
            rd.mass(basis, basis, rules, count, gb, geom, cfg.n_phys)
            rd.l2_dual(basis, rules, rng.standard_normal((count, cfg.Q)), gb, geom, cfg.n_phys)
    return time.perf_counter() - t0


def _port_step(name, count, seed):
    from oracle import oracle as orc

    cfg = syn.CONFIGS[name]
    basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
    rng = np.random.default_rng(seed)
    if name == "c2":
        dofs = rng.random((count, cfg.N))
        t0 = time.perf_counter()
        orc.reconstruct(basis, rules, dofs, [0, 1, 2])
        return time.perf_counter() - t0
    geom = _slab(cfg, 0, count)
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


class CpuArm:
    """The reference's CPU path on this host: `reference` = unmodified compiled interplib, one process per core;
    `port` = the C oracle restatement with OpenMP."""

    def __init__(self, name: str, single_thread: bool = False):
        from oracle import build_ref

        self.name = name
        self.kind = "reference" if build_ref.built() else "port"
        self.cores = 1 if single_thread else (os.cpu_count() or 1)
        self.pool = None
        if self.kind == "reference" and self.cores > 1:
            import multiprocessing as mp

            self.pool = mp.get_context("fork").Pool(self.cores)
        if self.kind == "port":
            from oracle import oracle as orc

            orc.build()
            if single_thread:
                os.environ["OMP_NUM_THREADS"] = "1"
            self.cores = 1 if single_thread else orc.threads()

    def run(self, count: int, seed: int = 0) -> float:
        """Process `count` elements; returns wall seconds."""
        t0 = time.perf_counter()
        if self.kind == "port":
            _port_step(self.name, count, seed)
        elif self.pool is None:
            _ref_worker((self.name, 0, count, seed))
        else:
            per = max(1, count // self.cores)
            jobs = [(self.name, i * per, per, seed + i) for i in range(self.cores)]
            self.pool.map(_ref_worker, jobs)
        return time.perf_counter() - t0

    def effective_count(self, count: int) -> int:
        if self.pool is None:
            return count
        return max(1, count // self.cores) * self.cores

    def calibrate(self, target_s: float) -> int:
        """Elements per step so that one step takes about target_s."""
        n = self.effective_count(max(self.cores * 4, 8))
        t = self.run(n)
        while t < 0.2 and n < (1 << 22):
            n *= 4
            t = self.run(n)
        return self.effective_count(max(self.cores, int(n * target_s / max(t, 1e-9))))

    def close(self):
        if self.pool:
            self.pool.close()
            self.pool.join()


def run_reference_arm(args, rank: int, world: int) -> int:
    if rank != 0:
        return 0
    name = args.workload
    cfg = syn.CONFIGS[name]
    arm = CpuArm(name)
    budget = 150.0  # seconds for warm-up + timed steps
    per_step = min(3.0, budget / max(args.steps + args.warmup, 1))
    count = arm.calibrate(per_step)
    for w in range(args.warmup):
        arm.run(count, seed=100 + w)
    t = [arm.run(count, seed=200 + k) for k in range(args.steps)]
    arm.close()
    total = float(np.sum(t))
    value = count * args.steps / total
    sample = (f"{count} elements per step of {cfg.name} ({cfg.description}); per-element calls through the "
              f"{'compiled reference Python API' if arm.kind == 'reference' else 'C oracle port'}")
    line = {
        "impl": "reference", "metric": "basis_grad_evals_per_s" if name == "c2" else "element_mass_matrices_per_s",
        "value": value, "unit": "elements/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{name}: {cfg.description}", "elements_per_step": count, "host_cores": arm.cores},
        "cpu_baseline": {"value": value, "unit": "elements/s", "cores": arm.cores, "kind": arm.kind, "sample": sample},
        "e2e": {"value": value, "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------ main
def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c2", choices=sorted(syn.CONFIGS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--elements", type=int, default=0, help="elements per rank (default: the configuration's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference_arm(args, rank, world)

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod

        torch.cuda.set_device(local_rank)
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist = dist_mod

    import interplib_b200 as ib

    numa = bind_to_gpu_numa_node(local_rank)
    ctx = ib.Context(local_rank)
    info = ctx.info()
    wl = Workload(args.workload, ctx, rank, args.elements or None)
    hbm_peak, peak_src = peaks()

    def barrier():
        ctx.synchronize()
        if dist is not None:
            import torch

            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        import torch

        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident timing -------------------------------------------------------------------------
    for _ in range(args.warmup):
        wl.step_device()
    barrier()
    sampler = ClockSampler()
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    barrier()
    # the W warm-up steps of a short step (C2: 0.3 ms) end long before the SM clocks have come up from idle, and the
    # sampler's start-up pause lets them fall again: keep stepping, untimed, until the GPU has been busy for 0.2 s
    t_warm, warm_run = time.perf_counter(), args.warmup
    while time.perf_counter() - t_warm < 0.2:
        for _ in range(8):
            wl.step_device()
        ctx.synchronize()
        warm_run += 8
    launches0 = ctx.stats()["kernel_launches"]
    barrier()
    ctx.timer_start(0)
    for _ in range(args.steps):
        wl.step_device()
    ctx.timer_stop(0)
    barrier()
    ms_total = max_over_ranks(ctx.timer_ms(0))
    launches = ctx.stats()["kernel_launches"] - launches0
    clocks = sampler.stop(set(range(world))) if rank == 0 else None
    ms_step = ms_total / args.steps
    value = wl.E * world / (ms_step * 1e-3)

    # average launch duration of the dominant kernel alone: CUDA events on the launching stream around back-to-back
    # launches of that kernel (its output is larger than L2, so consecutive launches do not hit in cache)
    n_dom = max(3, min(args.steps, 10))
    launches_before = ctx.stats()["kernel_launches"]
    ctx.timer_start(1)
    for _ in range(n_dom):
        wl.step_device(dominant_only=True)
    ctx.timer_stop(1)
    dom_launches = ctx.stats()["kernel_launches"] - launches_before
    kernel_ms = ctx.timer_ms(1) / max(dom_launches, 1)
    # a device-resident call larger than the context's scratch bound runs as several launches (C5: 1 863 680 elements each)
    el_per_launch = wl.E * n_dom / max(dom_launches, 1)
    achieved = wl.dominant_bytes_per_elem * el_per_launch / (kernel_ms * 1e-3) / 1e9
    traffic = NCU_TRAFFIC_BYTES.get(args.workload)
    if traffic is not None:
        traffic = int(round(traffic[0] * (el_per_launch / traffic[1])))
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                "traffic": traffic, "peak_source": peak_src, "kernel": wl.dominant,
                "algorithmic_bytes_per_element": wl.dominant_bytes_per_elem, "step_algorithmic_bytes_per_element": wl.bytes_per_elem,
                "step_frac": wl.bytes_per_elem * wl.E / (ms_step * 1e-3) / 1e9 / hbm_peak,
                "elements_per_launch": el_per_launch, "kernel_ms": kernel_ms,
                "fp64_flops_per_element": getattr(wl, "flops_per_elem", None)}

    # ---- end to end through the host-pointer API (pinned host buffers) -----------------------------------------
    e2e = None
    if not args.no_e2e:
        wl.e2e_prepare()
        n_e2e = args.e2e_steps or max(3, min(args.steps, 10))
        wl.step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            wl.step_e2e()
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": wl.E2 * world * n_e2e / dt, "unit": wl.unit, "h2d_bytes_per_step": wl.h2d_bytes,
               "d2h_bytes_per_step": wl.d2h_bytes, "steps": n_e2e, "ms_per_step": 1e3 * dt / n_e2e,
               "elements_per_step": wl.E2,
               "timing": "host wall clock around the synchronous C-ABI call, max over ranks", "cpu_affinity": numa}

    # ---- CPU baseline beside it (rank 0, N=1 only) -----------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        arm = CpuArm(args.workload, single_thread=True)
        count = arm.calibrate(12.0)
        t = arm.run(count, seed=7)
        arm.close()
        cpu = {"value": count / t, "unit": wl.unit, "cores": arm.cores, "kind": arm.kind,
               "sample": f"{count} elements of {args.workload} in {t:.1f} s, per-element calls, single thread (the reference "
                         "holds the GIL and has no parallel region)"}

    if rank == 0:
        cfg = wl.cfg
        line = {
            "metric": wl.metric, "value": value, "unit": wl.unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"{cfg.name}: {cfg.description}", "elements_per_gpu": wl.E, "basis": [list(b) for b in cfg.basis],
                       "rule": [list(r) for r in cfg.rules], "l2": "inputs+outputs larger than L2 "
                       f"({(wl.bytes_per_elem * wl.E) >> 20} MiB per step vs 126 MB L2)", "device": info["name"],
                       "sm_count": info["sm_count"], "parallelism": f"element-partitioned x{world}, no data-path collective",
                       "warmup_steps_run": warm_run},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches) * world, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
