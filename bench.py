#!/usr/bin/env python3
"""bench.py -- throughput of the interplib hot path on B200 (contract: DESIGN.md section 6).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c5|c4|c3|c1] [--impl ours|reference]

One "step" = one pass of the hot path over one batch of synthetic elements.  BASELINE.json's metric has two halves,
"basis+grad evals/s" and "element mass matrices/s"; the default run reports both in ONE JSON line:

* headline (`metric`, `value`, `roofline`, `e2e`, `cpu_baseline`): BASELINE config 2 -- 2-D quads, 2^20 elements per
  GPU, p=6, value + gradient at 8x8 Gauss points, FP64; weak scaling, no data-path collective;
* `mass`:   BASELINE config 5's element work on one 2^22-element slab per GPU -- 25x25 mass matrix + L2 right-hand side
  per element (`element_mass_matrices_per_s`), with its own `roofline`, `e2e` and `cpu_baseline`;
* `strong`: BASELINE config 5 as the north star states it -- 2^24 quads in total, block-partitioned over the N ranks,
  every rank runs the whole L2 projection of its elements (geometry + f in, primal DoFs out; mass matrices assembled
  and solved on the device) and the ranks then ALL-GATHER the [E,25] primal DoFs and ALL-REDUCE the projection norm
  over NCCL through the C ABI, all inside the CUDA-event region (`"scaling": "strong"` inside the record).

Elements are independent: under torchrun every rank owns a GPU; timing is on the device (CUDA events on the launching
stream), max over ranks.  Rank 0 prints the line.  torch is used for the launcher's barrier and the max over ranks
only; everything measured goes through libinterp_b200.so.

`--impl reference` times the reference's own CPU implementation (oracle/_ref: the unmodified interplib extension
driven per element through its Python API, one spawned worker process per host core; the C oracle port when the
compiled reference is absent) on a bounded sample of the same workloads and prints the same keys.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from interplib_b200 import synthetic as syn  # noqa: E402  (pure NumPy: does not load libinterp_b200.so)

# ncu-measured DRAM traffic of the dominant kernel (dram__bytes_read.sum + dram__bytes_write.sum of one launch from the
# `ncu --set full` captures summarised under profiles/*_full.txt) as (bytes per launch, elements in that launch); the
# bench line scales it to the elements of its own launch.  Slightly below the algorithmic bytes where the tail of the
# output is still dirty in the 126 MB L2 when the kernel ends.
NCU_TRAFFIC_BYTES = {
    "c2": (1970064432, 1 << 20, "profiles/r01e_c2_eval2d_full.txt (kernel unchanged since)"),
    "c5": (5249943208, 1 << 20, "profiles/r01c_c5_mass2d_lo_full.txt (kernel unchanged since)"),
    "c4": (43730439000 + 20404992, 1 << 16, "profiles/r02p_c4_mass2d_kron_staged_full.txt"),
    "c3": (43427470000 + 17022720, 1 << 18, "profiles/r02g_c3_mass2d_kron_full.txt"),
    "c1": None,
}
METRIC = {"c2": "basis_grad_evals_per_s"}
DEFAULT_ELEMENTS = {"c1": 1024, "c2": 1 << 20, "c3": 1 << 18, "c4": 1 << 16, "c5": 1 << 22}
# sum-factorised FP64 operation counts per element
FLOPS_PER_ELEMENT = {"c2": 2 * (2 * 7 * 7 * 8 + 3 * 7 * 8 * 8), "c4": 3.31e6, "c5": 10.9e3, "c3": 0.50e6}
# what the launched kernel really executes where that differs: the Kronecker-sum kernels of C3 / C4 spend one multiply and
# one fused multiply-add per matrix entry (3 moment-table terms folded into 2 operations) -- the nominal count
# above is what a sum-factorised assembly of the same matrices would cost and is reported as "effective" only
EXECUTED_FLOPS_PER_ELEMENT = {"c4": 2.0 * 289 * 289, "c3": 2.0 * 144 * 144}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def fp64_peak():
    """Measured FP64 peaks of this pool's B200 (tools/fp64_peak.cu -> profiles/r01_fp64_peak.json)."""
    p = ROOT / "profiles" / "r01_fp64_peak.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["fp64_fma_tflops"]), float(d["fp64_dmma_m8n8k4_tflops"]), "measured (profiles/r01_fp64_peak.json)"
    return 37.0, 37.0, "nominal"


def workload_config(name: str, elements_per_gpu: int) -> dict:
    """The `config` object: names the workload only, so the repo arm and the reference arm print the same one."""
    cfg = syn.CONFIGS[name]
    per_elem = {"c2": 8 * (cfg.N + 3 * cfg.Q), "c3": 8 * (8 + 144 * 144)}.get(name, 8 * (8 + cfg.Q + cfg.N * cfg.N + cfg.N))
    return {"workload": f"{cfg.name}: {cfg.description}", "elements_per_gpu": int(elements_per_gpu),
            "basis": [list(b) for b in cfg.basis], "rule": [list(r) for r in cfg.rules],
            "l2": f"inputs+outputs larger than L2 ({(per_elem * elements_per_gpu) >> 20} MiB per step vs 126 MB L2)",
            "parallelism": "element-partitioned, one rank per GPU, no data-path collective in this step"}


# ------------------------------------------------------------------------------------------------ workloads
class Workload:
    """Per-rank batch of one BASELINE configuration, device resident."""

    def __init__(self, name: str, ctx, rank: int, n_elements: int | None):
        self.name = name
        self.cfg = cfg = syn.CONFIGS[name]
        self.ctx = ctx
        self.basis, self.rules, self.gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
        E = n_elements or DEFAULT_ELEMENTS[name]
        self.E = E
        N, Q = cfg.N, cfg.Q
        rng = np.random.default_rng(2 + rank)
        self.blocks = None
        self.unit = "elements/s"
        self.metric = METRIC.get(name, "element_mass_matrices_per_s")
        self.flops_per_elem = FLOPS_PER_ELEMENT.get(name)
        if name == "c2":
            # basis + gradient eval: dofs[E][49] -> out[E][3][64]
            self.h_in = {"dofs": rng.random((E, N))}
            self.out_shapes = {"out": (E, 3, Q)}
            self.bytes_per_elem = 8 * (N + 3 * Q)
            self.dominant = "eval2d_kernel (K1: value + gradient)"
            self.dominant_bytes_per_elem = self.bytes_per_elem
            self.dominant_is_step = True
        elif name == "c3":
            # the determinant-weighted mass matrix of the 2-D 1-form basis (blocks M_ab = compute_mass_matrix(space_b, space_a,
            # smap), SURVEY 8c caveat): geometry[E][2][4] -> M[E][144][144] in ONE launch (ib200_kform_mass_mapped_batch)
            first = (rank * E) % max(cfg.n_elements - E + 1, 1)
            self.h_in = {"geom": syn.geometry(cfg, first=first, count=E)}
            self.kform_basis = [("lagrange-gauss-lobatto", 8)] * 2
            self.out_shapes = {"mass": (E, 144, 144)}
            self.bytes_per_elem = 8 * (8 + 144 * 144)
            self.dominant = "mass2d_kron_kernel<2,9> (K2-kron: the whole 144x144 1-form matrix per element, geometry fused, one launch)"
            self.dominant_bytes_per_elem = self.bytes_per_elem
            self.dominant_is_step = True
        else:
            # mass matrix (+ L2 right-hand side): geometry[E][2][4] (+ f[E][Q]) -> M[E][N][N] (+ b[E][N])
            first = (rank * E) % max(cfg.n_elements - E + 1, 1)
            if cfg.ndim == 2:
                geom = syn.geometry(cfg, first=first, count=E)
            else:
                geom = syn.geometry(cfg)[(np.arange(E)) % cfg.n_elements]
            self.h_in = {"geom": np.ascontiguousarray(geom), "fvals": rng.standard_normal((E, Q))}
            self.out_shapes = {"mass": (E, N, N), "rhs": (E, N)}
            self.bytes_per_elem = 8 * (geom[0].size + Q + N * N + N)
            fused = name == "c5"  # low order: geometry + mass + right-hand side are ONE kernel, which is then the step
            self.dominant = ({"c5": "mass2d_lo_kernel (K2-lo: geometry + mass matrix + L2 right-hand side fused)",
                              "c4": "mass2d_kron_kernel<1,17> (K2-kron, geometry fused)"}
                             .get(name, "mass2d_kernel (K2, geometry fused)") if cfg.ndim == 2 else "geometry + mass kernels (1-D generic)")
            self.dominant_bytes_per_elem = self.bytes_per_elem if fused else 8 * (geom[0].size + N * N)
            self.dominant_is_step = fused
        self.d_in = {k: ctx.alloc(v.nbytes).upload(v) for k, v in self.h_in.items()}
        self.d_out = {k: ctx.alloc(int(np.prod(s)) * 8) for k, s in self.out_shapes.items()}
        self.pinned = []

    def free(self) -> None:
        for b in list(self.d_in.values()) + list(self.d_out.values()):
            b.free()
        for p in self.pinned:
            p.free()
        self.h_in = self.d_in = self.d_out = None

    @property
    def E2(self) -> int:
        """Elements of one end-to-end step: the host-pointer path streams its arrays through PCIe in 16 MB chunks, so its
        rate does not depend on the batch; C3-C5 use a slab of the device-resident batch to bound pinned host memory."""
        return min(self.E, {"c3": 1 << 15, "c4": 1 << 13, "c5": 1 << 20}.get(self.name, self.E))

    def step_device(self, dominant_only: bool = False) -> None:
        from interplib_b200 import DEVICE

        c = self.ctx
        if self.name == "c2":
            c.reconstruct(self.basis, self.rules, self.d_in["dofs"].ptr, [0, 1, 2], out=self.d_out["out"].ptr, where=DEVICE,
                          n_elements=self.E)
        elif self.name == "c3":
            c.kform_mass_mapped(self.kform_basis, 1, self.rules, self.gb, self.d_in["geom"].ptr, 2, 144, out=self.d_out["mass"].ptr,
                                where=DEVICE, n_elements=self.E)
        else:
            if dominant_only and not self.dominant_is_step:
                c.mass_matrix_mapped(self.basis, self.basis, self.rules, self.gb, self.d_in["geom"].ptr, self.cfg.n_phys,
                                     out=self.d_out["mass"].ptr, where=DEVICE, n_elements=self.E)
                return
            # mass matrix + L2 right-hand side of the same elements: one fused kernel at low order (K2-lo + K3),
            # two launches otherwise (ib200_l2_system_mapped_batch)
            c.l2_system_mapped(self.basis, self.rules, self.gb, self.d_in["geom"].ptr, self.cfg.n_phys, self.d_in["fvals"].ptr,
                               mass=self.d_out["mass"].ptr, rhs=self.d_out["rhs"].ptr, where=DEVICE, n_elements=self.E)

    # ---- end to end through the public host-pointer API (pinned host buffers, H2D + D2H inside the call) ------------
    def e2e_prepare(self, chain: bool = False) -> None:
        c = self.ctx
        n = self.E2
        self.p_in = {k: c.pinned((n,) + v.shape[1:]) for k, v in self.h_in.items()}
        for k, v in self.h_in.items():
            self.p_in[k].array[...] = v[:n]
        shapes = {"primal": (n, self.cfg.N)} if chain else {k: (n,) + tuple(s[1:]) for k, s in self.out_shapes.items()}
        self.p_out = {k: c.pinned(s) for k, s in shapes.items()}
        self.pinned += list(self.p_in.values()) + list(self.p_out.values())

    def e2e_release(self) -> None:
        for p in self.pinned:
            p.free()
        self.pinned = []

    def step_e2e(self) -> float:
        c = self.ctx
        if self.name == "c2":
            c.reconstruct(self.basis, self.rules, self.p_in["dofs"].array, [0, 1, 2], out=self.p_out["out"].array)
            return float(self.p_out["out"].array[-1, 0, 0])
        if self.name == "c3":
            c.kform_mass_mapped(self.kform_basis, 1, self.rules, self.gb, self.p_in["geom"].array, 2, 144, out=self.p_out["mass"].array)
            return float(self.p_out["mass"].array[-1, 0, 0])
        c.l2_system_mapped(self.basis, self.rules, self.gb, self.p_in["geom"].array, self.cfg.n_phys, self.p_in["fvals"].array,
                           mass=self.p_out["mass"].array, rhs=self.p_out["rhs"].array)
        return float(self.p_out["rhs"].array[-1, 0])

    def step_e2e_chain(self) -> float:
        """projection_l2_primal of the slab through the host-pointer C ABI: geometry + f in, primal DoFs out; the mass
        matrices are assembled, factorised and consumed on the device (ib200_l2_project_mapped_batch)."""
        self.ctx.l2_project_mapped(self.basis, self.rules, self.gb, self.p_in["geom"].array, self.cfg.n_phys,
                                   self.p_in["fvals"].array, out=self.p_out["primal"].array)
        return float(self.p_out["primal"].array[-1, 0])

    def h2d_bytes(self) -> int:
        return int(sum(v.nbytes // self.E * self.E2 for v in self.h_in.values()))

    def d2h_bytes(self, chain: bool = False) -> int:
        if chain:
            return 8 * self.cfg.N * self.E2
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
        sm, smax, power, reasons = [], [], [], set()
        for line in Path(self.path).read_text().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9 or not f[0].isdigit() or int(f[0]) not in gpu_indices:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            try:
                power.append(float(f[3]))
            except ValueError:
                pass
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "sm_min_mhz": min(sm) if sm else None, "power_w": float(np.median(power)) if power else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ reference arm
def _reference_record(name: str, steps: int, warmup: int, budget_s: float, elements_cfg: int) -> dict:
    """Time the reference's CPU path on all host cores for one workload; every step is a bounded sample."""
    from oracle.cpu_arm import CpuArm

    cfg = syn.CONFIGS[name]
    arm = CpuArm(name)
    per_step = min(4.0, budget_s / max(steps + warmup, 1))
    count = arm.calibrate(per_step, cap=elements_cfg)
    for w in range(warmup):
        arm.run(count, seed=100 + w)
    t = [arm.run(count, seed=200 + k) for k in range(steps)]
    arm.close()
    total = float(np.sum(t))
    value = count * steps / total
    sample = (f"{count} elements per step of {cfg.name} ({cfg.description}); per-element calls through the "
              f"{'compiled reference Python API, one spawned worker process per core' if arm.kind == 'reference' else 'C oracle port (OpenMP)'}")
    return {"metric": METRIC.get(name, "element_mass_matrices_per_s"), "value": value, "unit": "elements/s",
            "ms_per_step": 1e3 * total / steps, "steps": steps, "warmup": warmup,
            "config": workload_config(name, elements_cfg),
            "cpu_baseline": {"value": value, "unit": "elements/s", "cores": arm.cores, "kind": arm.kind, "sample": sample},
            "e2e": {"value": value, "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def run_reference_arm(args, rank: int, world: int) -> int:
    if rank != 0:
        return 0
    name = args.workload
    head = _reference_record(name, args.steps, args.warmup, 110.0, args.elements or DEFAULT_ELEMENTS[name])
    line = {"impl": "reference", "metric": head["metric"], "value": head["value"], "unit": head["unit"], "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": head["config"],
            "cpu_baseline": head["cpu_baseline"], "e2e": head["e2e"], "gpu_launches": 0}
    if name == "c2" and not args.no_mass:
        # the other half of the metric: config 5's per-element mass matrix + L2 right-hand side
        line["mass"] = _reference_record("c5", 3, 1, 40.0, DEFAULT_ELEMENTS["c5"])
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------ our arm
class Runner:
    def __init__(self, args):
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        if self.world > 1:
            import torch
            import torch.distributed as dist_mod

            torch.cuda.set_device(self.local_rank)
            dist_mod.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
            self.dist = dist_mod
        import interplib_b200 as ib

        self.numa = bind_to_gpu_numa_node(self.local_rank)
        self.ctx = ib.Context(self.local_rank)
        self.info = self.ctx.info()
        self.hbm_peak, self.peak_src = peaks()

    def barrier(self):
        self.ctx.synchronize()
        if self.dist is not None:
            import torch

            self.dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.dist is None:
            return x
        import torch

        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{self.local_rank}")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def keep_busy(self, step, seconds: float = 0.2) -> int:
        """A short step's W warm-ups end long before the SM clocks have come up from idle: keep stepping, untimed, until
        the GPU has been busy for `seconds` right before the timed region.  Returns the extra steps run."""
        t0, n = time.perf_counter(), 0
        while time.perf_counter() - t0 < seconds:
            for _ in range(8):
                step()
            self.ctx.synchronize()
            n += 8
        return n

    # ---- one workload: device-resident value + roofline + e2e + CPU baseline ------------------------------------------
    def measure(self, name: str, steps: int, warmup: int, elements: int | None, sample_clocks: bool, chain_e2e: bool,
                do_e2e: bool = True) -> dict:
        args, ctx, rank, world = self.args, self.ctx, self.rank, self.world
        wl = Workload(name, ctx, rank, elements)
        for _ in range(warmup):
            wl.step_device()
        self.barrier()
        sampler = ClockSampler()
        if rank == 0 and sample_clocks:
            sampler.start()
            time.sleep(0.25)
        self.barrier()
        warm_run = warmup + self.keep_busy(wl.step_device)
        launches0 = ctx.stats()["kernel_launches"]
        self.barrier()
        ctx.timer_start(0)
        for _ in range(steps):
            wl.step_device()
        ctx.timer_stop(0)
        self.barrier()
        ms_total_local = ctx.timer_ms(0)
        ms_total = self.max_over_ranks(ms_total_local)
        launches = ctx.stats()["kernel_launches"] - launches0
        clocks = sampler.stop(set(range(world))) if (rank == 0 and sample_clocks) else None
        ms_step = ms_total / steps
        value = wl.E * world / (ms_step * 1e-3)

        # average launch duration of the dominant kernel: when the step IS that kernel (C2: one eval2d launch; C5: the fused
        # geometry + mass + right-hand-side launches) it comes from the step's own timed region, otherwise from a second
        # CUDA-event region of the same length around that kernel alone (outputs larger than L2: no cache hits between launches)
        if wl.dominant_is_step:
            dom_launches, dom_ms_total, n_dom = launches, ms_total_local, steps
        else:
            n_dom = steps
            before = ctx.stats()["kernel_launches"]
            ctx.timer_start(1)
            for _ in range(n_dom):
                wl.step_device(dominant_only=True)
            ctx.timer_stop(1)
            dom_launches = ctx.stats()["kernel_launches"] - before
            dom_ms_total = ctx.timer_ms(1)
        kernel_ms = dom_ms_total / max(dom_launches, 1)
        el_per_launch = wl.E * n_dom / max(dom_launches, 1)
        achieved = wl.dominant_bytes_per_elem * el_per_launch / (kernel_ms * 1e-3) / 1e9
        traffic = traffic_src = NCU_TRAFFIC_BYTES.get(name)
        if traffic is not None:
            traffic_src = f"ncu --set full capture {traffic[2]}, {traffic[0]} B per {traffic[1]} elements, scaled to this launch"
            traffic = int(round(traffic[0] * (el_per_launch / traffic[1])))
        roofline = {"bound": "hbm", "achieved": achieved, "peak": self.hbm_peak, "unit": "GB/s", "frac": achieved / self.hbm_peak,
                    "traffic": traffic, "traffic_source": traffic_src, "peak_source": self.peak_src, "kernel": wl.dominant,
                    "algorithmic_bytes_per_element": wl.dominant_bytes_per_elem,
                    "step_algorithmic_bytes_per_element": wl.bytes_per_elem,
                    "step_frac": wl.bytes_per_elem * wl.E / (ms_total_local / steps * 1e-3) / 1e9 / self.hbm_peak,
                    "elements_per_launch": el_per_launch, "kernel_ms": kernel_ms,
                    "kernel_ms_source": "the step's own CUDA-event region" if wl.dominant_is_step else "separate CUDA-event region, dominant kernel alone",
                    "fp64_effective_flops_per_element": wl.flops_per_elem}
        if wl.flops_per_elem:
            fma, dmma, src = fp64_peak()
            step_s = ms_total_local / steps * 1e-3
            tf = wl.flops_per_elem * wl.E / step_s / 1e12
            executed = EXECUTED_FLOPS_PER_ELEMENT.get(name, wl.flops_per_elem)
            tf_x = executed * wl.E / step_s / 1e12
            # "effective" = the SURVEY's sum-factorised operation count over the step time (what an assembly by quadrature
            # would have needed); "executed" = what this step's kernels really issue -- equal except for the Kronecker kernels
            roofline.update({"fp64_effective_tflops": tf, "fp64_effective_frac": tf / fma,
                             "fp64_executed_flops_per_element": executed, "fp64_executed_tflops": tf_x,
                             "fp64_executed_frac": tf_x / fma, "fp64_peak_tflops": fma, "fp64_peak_source": src})

        # ---- end to end through the host-pointer API (pinned host buffers) ----------------------------------------------
        def time_e2e(step, n):
            step()
            self.barrier()
            t0 = time.perf_counter()
            for _ in range(n):
                step()
            self.barrier()
            return self.max_over_ranks(time.perf_counter() - t0)

        e2e = e2e_full = projection_e2e = None
        if not args.no_e2e and do_e2e:
            n_e2e = args.e2e_steps or max(3, min(steps, 10))
            note = {"steps": n_e2e, "elements_per_step": wl.E2, "cpu_affinity": self.numa,
                    "timing": "host wall clock around the synchronous C-ABI call with pinned host buffers, max over ranks"}
            wl.e2e_prepare(chain=False)
            dt = time_e2e(wl.step_e2e, n_e2e)
            e2e_full = {"value": wl.E2 * world * n_e2e / dt, "unit": wl.unit, "h2d_bytes_per_step": wl.h2d_bytes(),
                        "d2h_bytes_per_step": wl.d2h_bytes(), "ms_per_step": 1e3 * dt / n_e2e, **note,
                        "path": "every output array of the step travels to the host"}
            wl.e2e_release()
            e2e = e2e_full
            if chain_e2e:
                wl.e2e_prepare(chain=True)
                dt = time_e2e(wl.step_e2e_chain, n_e2e)
                projection_e2e = {
                    "metric": "l2_projections_per_s", "value": wl.E2 * world * n_e2e / dt, "unit": wl.unit,
                    "h2d_bytes_per_step": wl.h2d_bytes(), "d2h_bytes_per_step": wl.d2h_bytes(chain=True),
                    "ms_per_step": 1e3 * dt / n_e2e, **note,
                    "path": "whole L2 projection in one call (ib200_l2_project_mapped_batch): geometry + f in, primal DoFs out. "
                            "NOT a mass-matrix number: on these bilinear quads / straight segments the default solver is the "
                            "fast diagonalisation (1-D projector), so no mass matrix is formed at all"}
                wl.e2e_release()

        # ---- CPU baseline beside it (rank 0, N=1 only) ------------------------------------------------------------------
        cpu = None
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            from oracle.cpu_arm import CpuArm

            arm = CpuArm(name, single_thread=True)
            count = arm.calibrate(10.0)
            t = arm.run(count, seed=7)
            arm.close()
            cpu = {"value": count / t, "unit": wl.unit, "cores": arm.cores, "kind": arm.kind,
                   "sample": f"{count} elements of {name} in {t:.1f} s, per-element calls, single thread (the reference "
                             "holds the GIL and has no parallel region)"}
        rec = {"metric": wl.metric, "value": value, "unit": wl.unit, "steps": steps, "warmup": warmup, "ms_per_step": ms_step,
               "config": workload_config(name, wl.E), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
               "gpu_launches": int(launches) * world, "clocks": clocks,
               "run": {"device": self.info["name"], "sm_count": self.info["sm_count"], "warmup_steps_run": warm_run}}
        if projection_e2e is not None:
            rec["projection_e2e"] = projection_e2e
        wl.free()
        return rec

    # ---- BASELINE config 5 as a strong-scaling job with the NCCL gather / reduce inside the timed region -----------------
    def measure_strong(self, total_elements: int, steps: int, warmup: int) -> dict:
        from interplib_b200 import DEVICE
        from interplib_b200.partition import ElementPartition, NcclComm

        ctx, rank, world = self.ctx, self.rank, self.world
        cfg = syn.CONFIGS["c5"]
        basis, rules, gb = list(cfg.basis), list(cfg.rules), list(cfg.geom_basis)
        N, Q = cfg.N, cfg.Q
        comm = NcclComm.from_environment(ctx)
        part = ElementPartition.from_comm(total_elements, comm)
        base_n = min(1 << 20, total_elements)
        f_base = np.random.default_rng(99).standard_normal((base_n, Q))  # f[e] = f_base[e mod 2^20]: the same on every rank

        def upload_inputs(first: int, count: int):
            d_g = ctx.alloc(max(count, 1) * 64)
            d_f = ctx.alloc(max(count, 1) * Q * 8)
            done = 0
            while done < count:  # geometry of the real 4096 x 4096 jittered mesh, generated in slabs
                n = min(1 << 21, count - done)
                g = syn.geometry(cfg, first=first + done, count=n)
                ctx_copy_h2d(ctx, int(d_g) + done * 64, g)
                done += n
            done = 0
            while done < count:
                lo = (first + done) % base_n
                n = min(base_n - lo, count - done)
                ctx_copy_h2d(ctx, int(d_f) + done * Q * 8, f_base[lo : lo + n])
                done += n
            return d_g, d_f

        d_all = ctx.alloc(total_elements * N * 8)      # the gathered [E][25] primal DoFs, every rank
        # gather destination registered with the communicator: slabs travel as copy-engine pushes into the peers' IPC-mapped
        # buffers (NVLink) instead of NCCL kernels; False -> NCCL broadcasts ($IB200_NO_P2P_GATHER=1 forces that)
        p2p = comm.register_gather(d_all, total_elements * N * 8)
        n_sub = 4                                       # sub-slabs per rank: the gather of one overlaps the assembly of the same and the next
        sub = -(-part.max_count // n_sub)
        d_norm = ctx.alloc(8 * n_sub)

        n_gather = max(1, int(os.environ.get("IB200_BENCH_GATHER_PARTS", "1")))  # pieces the DoF gather is issued in

        def job(d_g, d_f, d_b, d_M, first: int, count: int, collectives: bool, compute: bool = True, sub: int = sub,
                gather: bool = True):
            """Mass matrices + L2 projection of `count` elements starting at global element `first` (this rank's block, or all
            of them for the single-rank reference): primal DoFs of the block first (fast diagonalisation), their all-gather
            goes to the communicator's stream, the (M, b) assembly and the norms run meanwhile in n_sub sub-slabs."""
            gsub = -(-part.max_count // n_gather)
            for k in range(n_gather):
                lo, hi = min(count, k * gsub), min(count, (k + 1) * gsub)
                if hi > lo and compute:
                    ctx.l2_project_mapped(basis, rules, gb, int(d_g) + lo * 64, cfg.n_phys, int(d_f) + lo * Q * 8,
                                          out=int(d_all) + (first + lo) * N * 8, where=DEVICE, n_elements=hi - lo)
                if collectives and gather:
                    counts, offsets = part.slab_layout(N, k * gsub, gsub)
                    comm.allgatherv(int(d_all) + (part.first + min(part.count, k * gsub)) * N * 8, d_all, counts, offsets, overlap=True)
            for k in range(n_sub):
                lo, hi = min(count, k * sub), min(count, (k + 1) * sub)
                if hi > lo and compute:
                    g, f = int(d_g) + lo * 64, int(d_f) + lo * Q * 8
                    u = int(d_all) + (first + lo) * N * 8
                    b = int(d_b) + lo * N * 8
                    ctx.l2_system_mapped(basis, rules, gb, g, cfg.n_phys, f, mass=int(d_M) + lo * N * N * 8, rhs=b, where=DEVICE,
                                         n_elements=hi - lo)
                    ctx.dot(b, u, out=int(d_norm) + 8 * k, where=DEVICE, n=(hi - lo) * N)
            if collectives:
                comm.allreduce_sums(int(d_norm), n_sub)
                comm.join()

        def timed(fn, k_steps):
            self.barrier()
            ctx.timer_start(2)
            for _ in range(k_steps):
                fn()
            ctx.timer_stop(2)
            self.barrier()
            return self.max_over_ranks(ctx.timer_ms(2)) / k_steps

        d_g, d_f = upload_inputs(part.first, part.count)
        d_b = ctx.alloc(max(part.count, 1) * N * 8)
        d_M = ctx.alloc(max(part.count, 1) * N * N * 8)   # this rank's element mass matrices: stay sharded, never gathered
        run_all = lambda: job(d_g, d_f, d_b, d_M, part.first, part.count, True)  # noqa: E731
        for _ in range(max(warmup, 3)):
            run_all()
        sampler = ClockSampler()
        if rank == 0:
            sampler.start()
        # ~1 s of the same job before the timed legs, so that clocks and board power have reached their loaded state (the job
        # runs into the power cap: 1965 -> ~1790 MHz); the count comes from a max-over-ranks timing, hence equal on every rank
        n_pre = int(min(400, max(20, 1000.0 / max(timed(run_all, 3), 1e-3))))
        for _ in range(n_pre):
            run_all()
        launches0 = ctx.stats()["kernel_launches"]
        ms_step = timed(run_all, steps)
        launches = ctx.stats()["kernel_launches"] - launches0
        ms_compute = timed(lambda: job(d_g, d_f, d_b, d_M, part.first, part.count, False), steps)
        ms_coll = timed(lambda: job(d_g, d_f, d_b, d_M, part.first, part.count, True, compute=False), steps)
        # the DoF gather is "only on request" -- the same job with the norms all-reduced and the DoFs left sharded
        ms_reduce_only = timed(lambda: job(d_g, d_f, d_b, d_M, part.first, part.count, True, gather=False), steps)
        clocks = sampler.stop(set(range(world))) if rank == 0 else None
        comm.check()

        # ---- parity of the collective, on a step that cannot borrow from earlier ones: every PEER slab of d_all is filled
        # with NaN first (barrier on both sides, so no push of the checked step can precede the fill), then one whole step
        # runs, then the head of EVERY rank's slab -- recomputed by this rank alone -- must equal the rows the all-gather
        # delivered, bit for bit.  A push that never arrived or arrived after the join leaves NaN behind.
        # (GPUTEST boxes have one GPU: this is the driver-visible check of the collectives.)
        from interplib_b200._lib import check as _check, lib as _lib

        ctx.synchronize()
        self.barrier()
        own_lo, own_hi = part.first * N * 8, (part.first + part.count) * N * 8
        if own_lo:
            _check(_lib.ib200_device_memset(ctx.handle, int(d_all), 0xFF, own_lo))
        if total_elements * N * 8 > own_hi:
            _check(_lib.ib200_device_memset(ctx.handle, int(d_all) + own_hi, 0xFF, total_elements * N * 8 - own_hi))
        ctx.synchronize()
        self.barrier()
        run_all()
        ctx.synchronize()
        proj_sq = float(np.sum(d_norm.download((n_sub,))))  # of a whole step (the collective-only pass re-reduces in place)
        parity_local, checked = True, 0
        for r in range(world):
            chk_n = min(4096 if r == world - 1 else 512, part.counts[r])
            if chk_n == 0:
                continue
            chk_first = part.offsets[r]
            c_g, c_f = upload_inputs(chk_first, chk_n)
            c_u = ctx.alloc(chk_n * N * 8)
            ctx.l2_project_mapped(basis, rules, gb, int(c_g), cfg.n_phys, int(c_f), out=int(c_u), where=DEVICE, n_elements=chk_n)
            ctx.synchronize()
            mine = c_u.download((chk_n, N))
            got = d_all.download((chk_n, N), offset_bytes=chk_first * N * 8)
            parity_local = parity_local and bool(np.array_equal(mine, got)) and bool(np.isfinite(got).all())
            checked += chk_n
            for b in (c_g, c_f, c_u):
                b.free()
        # the LAST row of every peer slab as well: the tail of a push is what a late join would miss
        for r in range(world):
            if part.counts[r]:
                last = d_all.download((1, N), offset_bytes=(part.offsets[r] + part.counts[r] - 1) * N * 8)
                parity_local = parity_local and bool(np.isfinite(last).all())
        parity = self.max_over_ranks(0.0 if parity_local else 1.0) == 0.0
        for b in (d_g, d_f, d_b, d_M):
            b.free()

        # ---- the same 2^24 elements on ONE GPU (rank 0 alone, no collective): the denominator of the strong-scaling efficiency
        ms_single = clocks_single = None
        if world > 1:
            if rank == 0:
                s_g, s_f = upload_inputs(0, total_elements)
                s_b = ctx.alloc(total_elements * N * 8)
                s_M = ctx.alloc(total_elements * N * N * 8)
                one = lambda: job(s_g, s_f, s_b, s_M, 0, total_elements, False, sub=-(-total_elements // n_sub))  # noqa: E731
                sampler1 = ClockSampler()
                sampler1.start()
                one()
                ctx.synchronize()
                t0 = time.perf_counter()
                while time.perf_counter() - t0 < 1.0:   # the same ~1 s under load as the multi-rank legs get
                    one()
                    ctx.synchronize()
                k1 = max(3, min(steps, 5))
                ctx.timer_start(3)
                for _ in range(k1):
                    one()
                ctx.timer_stop(3)
                ms_single = ctx.timer_ms(3) / k1
                clocks_single = sampler1.stop({self.local_rank})
                for b in (s_g, s_f, s_b, s_M):
                    b.free()
            self.barrier()
            ms_single = self.max_over_ranks(ms_single or 0.0)
        else:
            ms_single = ms_compute
        nccl = comm.info()["nccl_version"]
        comm.unregister_gather()
        comm.close()
        d_all.free()
        d_norm.free()
        gather_bytes_in = (total_elements - part.max_count) * N * 8  # bytes every GPU receives
        link_bound_ms = gather_bytes_in / 900e9 * 1e3
        # algorithmic bytes of the whole job per element: geometry + f in, M + b + u out (SURVEY 8d C5 row + the primal DoFs)
        job_bytes = 8 * (8 + Q + N * N + 2 * N)
        rec = {
            "metric": "mass_matrix_and_l2_projection_per_s", "unit": "elements/s", "scaling": "strong",
            "value": total_elements / (ms_step * 1e-3), "ms_per_step": ms_step, "steps": steps,
            "config": {"workload": f"c5: {cfg.description}", "total_elements": total_elements, "elements_per_gpu": part.max_count,
                       "basis": [list(b) for b in cfg.basis], "rule": [list(r) for r in cfg.rules],
                       "step": "per rank, for its block of the 16M quads: primal DoFs u (ib200_l2_project_mapped_batch, fast "
                               "diagonalisation), element mass matrices M + right-hand sides b (ib200_l2_system_mapped_batch; M stays "
                               "resident and sharded, 5 000 B per element), norm b.u (ib200_dot); then NCCL all-gather of the [E,25] primal "
                               "DoFs to every rank and all-reduce of the norms, issued through the C ABI, all inside the CUDA-event region",
                       "pipeline": f"primal DoFs of the rank's block first; their all-gather ({n_gather} part(s)) runs on the communicator's "
                                   f"(high-priority) stream while the mass matrices and right-hand sides are assembled in {n_sub} sub-slabs",
                       "parallelism": f"element block partition x{world}, NCCL {nccl} all-gather + all-reduce",
                       "gather_path": ("copy-engine pushes into IPC-mapped peer buffers over NVLink (ib200_comm_register_gather), arrival "
                                       "published by a one-double NCCL all-reduce") if p2p else "NCCL broadcasts grouped per sub-slab"},
            "compute_only_ms": ms_compute, "collective_only_ms": ms_coll, "exposed_collective_ms": ms_step - ms_compute,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_element": job_bytes,
                         "achieved": job_bytes * part.max_count / (ms_compute * 1e-3) / 1e9, "peak": self.hbm_peak, "unit": "GB/s",
                         "frac": job_bytes * part.max_count / (ms_compute * 1e-3) / 1e9 / self.hbm_peak,
                         "note": "compute-only pass of one rank (no collective): three launches per sub-slab"},
            "collective": {"allgather_bytes_received_per_gpu": gather_bytes_in,
                           "allgather_gbs_per_gpu": (gather_bytes_in / (ms_coll * 1e-3) / 1e9) if ms_coll > 0 and world > 1 else None,
                           "link_peak_gbs_per_direction": 900.0,
                           "link_bound_ms": link_bound_ms},
            "single_gpu_ms": ms_single, "speedup_vs_single_gpu": ms_single / ms_step,
            "efficiency": ms_single / ms_step / world,
            "compute_efficiency": ms_single / ms_compute / world,
            # what NVLink's 900 GB/s into every GPU allows for this job: the gathered DoFs (3.4 GB on every rank) cannot arrive
            # faster than link_bound_ms, however well the transfer hides behind the kernels
            "link_bound_efficiency": ms_single / max(ms_single / world, link_bound_ms) / world,
            "reduce_only": {"ms_per_step": ms_reduce_only, "value": total_elements / (ms_reduce_only * 1e-3),
                            "speedup_vs_single_gpu": ms_single / ms_reduce_only, "efficiency": ms_single / ms_reduce_only / world,
                            "step": "the same job with the norms all-reduced (NCCL) and the primal DoFs left sharded on their ranks "
                                    "(SURVEY 8e: the DoF gather is only on request)"},
            "clocks": clocks, "clocks_single_gpu_leg": clocks_single,
            "collective_parity": parity, "collective_parity_check": f"peer slabs NaN-filled before the checked step; {checked} rows over all {world} rank slabs recomputed locally, bit-equal; last row of every slab finite", "projection_norm_sq": proj_sq, "gpu_launches": int(launches) * world,
        }
        return rec


def ctx_copy_h2d(ctx, dptr: int, host: np.ndarray) -> None:
    import ctypes as C

    from interplib_b200._lib import check, lib

    host = np.ascontiguousarray(host)
    check(lib.ib200_copy_h2d(ctx.handle, C.c_void_p(dptr), host.ctypes.data_as(C.c_void_p), host.nbytes))


# ------------------------------------------------------------------------------------------------ main
def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c2", choices=sorted(syn.CONFIGS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--elements", type=int, default=0, help="elements per rank (default: the configuration's)")
    ap.add_argument("--strong-elements", type=int, default=1 << 24, help="total elements of the strong-scaling record")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-mass", action="store_true", help="skip the config-5 `mass` record of the default run")
    ap.add_argument("--no-others", action="store_true", help="skip the config-1/3/4 records of the default run")
    ap.add_argument("--no-strong", action="store_true", help="skip the config-5 strong-scaling record of the default run")
    ap.add_argument("--e2e-steps", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        return run_reference_arm(args, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))

    r = Runner(args)
    name = args.workload
    head = r.measure(name, args.steps, args.warmup, args.elements or None, sample_clocks=True, chain_e2e=name in ("c5", "c4", "c1"))
    mass = strong = None
    others = {}
    if name == "c2":
        if not args.no_others:
            # BASELINE configs 1, 3, 4: device-resident value + roofline + single-thread CPU baseline (no e2e: 44 GB of
            # matrices per step over PCIe says nothing the C5 `mass.e2e` does not say already)
            for other in ("c1", "c3", "c4"):
                others[other] = r.measure(other, max(5, min(args.steps, 10)), 3, None, sample_clocks=True, chain_e2e=False,
                                          do_e2e=False)
        if not args.no_mass:
            mass = r.measure("c5", max(5, min(args.steps, 20)), 3, None, sample_clocks=True, chain_e2e=True)
        if not args.no_strong:
            strong = r.measure_strong(args.strong_elements, max(5, min(args.steps, 10)), 3)
    if r.rank == 0:
        line = {
            "metric": head["metric"], "value": head["value"], "unit": head["unit"], "n_gpus": r.world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": head["config"], "roofline": head["roofline"],
            "cpu_baseline": head["cpu_baseline"], "e2e": head["e2e"],
            "gpu_launches": head["gpu_launches"] + (mass["gpu_launches"] if mass else 0) + (strong["gpu_launches"] if strong else 0)
            + sum(o["gpu_launches"] for o in others.values()),
            "clocks": head["clocks"], "run": head["run"],
        }
        if "projection_e2e" in head:
            line["projection_e2e"] = head["projection_e2e"]
        for other, rec in others.items():
            rec["us_per_step"] = 1e3 * rec["ms_per_step"]
            line[other] = rec
        if mass is not None:
            line["mass"] = mass
        if strong is not None:
            line["strong"] = strong
        print(json.dumps(line), flush=True)
    if r.dist is not None:
        r.dist.barrier()
        r.dist.destroy_process_group()
    r.ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
