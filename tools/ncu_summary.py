#!/usr/bin/env python3
"""Turn an `ncu --set full` report (gpurun_out/*.ncu-rep) into the small text summary committed under profiles/.

    python tools/ncu_summary.py gpurun_out/prof_c2.ncu-rep > profiles/r01_c2_eval2d_full.txt
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__block_size",
    "launch__grid_size", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "launch__waves_per_multiprocessor", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
    "smsp__inst_executed.sum", "sm__sass_thread_inst_executed_op_dfma_pred_on.sum",
    "sm__sass_thread_inst_executed_op_dmul_pred_on.sum", "sm__sass_thread_inst_executed_op_dadd_pred_on.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
    "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "lts__t_sectors_op_write.sum", "lts__t_sectors_op_read.sum",
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(f"# kernel: {d.get('Kernel Name')}\n# grid {d.get('Grid Size')} block {d.get('Block Size')}  (source: {rep})")
        for k in KEYS:
            if k in d:
                print(f"{k:72s} {d[k]:>18s} {units[hdr.index(k)]}")
        stalls = [(float(d[k] or 0), k) for k in hdr if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("_per_issue_active.ratio") or
                  (k.startswith("smsp__average_warp_latency_issue_stalled") and k.endswith(".ratio"))]
        if not stalls:
            stalls = [(float(d[k] or 0), k) for k in hdr if "issue_stalled" in k and k.endswith(".pct")]
        print("# warp stall reasons (largest first)")
        for v, k in sorted(stalls, reverse=True)[:8]:
            print(f"{k:72s} {v:18.3f}")
        try:
            rd, wr = float(d["dram__bytes_read.sum"]), float(d["dram__bytes_write.sum"])
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            rd *= scale[units[hdr.index("dram__bytes_read.sum")]]
            wr *= scale[units[hdr.index("dram__bytes_write.sum")]]
            t = float(d["gpu__time_duration.sum"]) * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1}[units[hdr.index("gpu__time_duration.sum")]]
            print(f"# dram traffic per launch = {rd + wr:.0f} bytes; under-ncu duration {t * 1e6:.1f} us -> {(rd + wr) / t / 1e9:.0f} GB/s DRAM")
        except (KeyError, ValueError):
            pass
        print()


if __name__ == "__main__":
    main()
