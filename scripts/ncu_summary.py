#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into the handful of metrics this path is judged on.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [kernel-substring]"""
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__inst_issued.sum",
    "smsp__inst_executed_pipe_fp64.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__thread_inst_executed_per_inst_executed.pct",
    "smsp__sass_average_branch_targets_threads_uniform.pct",
    "sm__sass_branch_targets_threads_divergent.sum", "sm__sass_branch_targets.sum",
    "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum",
]
STALL = "smsp__average_warps_issue_stalled_"
STALL2 = "smsp__average_warp_latency_issue_stalled_"


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = d.get("Kernel Name", "?")
        if len(sys.argv) > 2 and sys.argv[2] not in name:
            continue
        print(f"## {name}  (id {d.get('ID')})")
        u = dict(zip(hdr, units))
        for k in KEYS:
            if k in d:
                print(f"  {k:75s} {d[k]:>18s} {u[k]}")
        stalls = sorted(((float(d[k].replace(',', '')), k) for k in d if (k.startswith(STALL) or k.startswith(STALL2)) and k.endswith("_per_warp_active.pct") is False and d[k] not in ("", "n/a")), reverse=True)
        for v, k in stalls[:10]:
            print(f"  {k:75s} {v:18.3f} {u[k]}")


if __name__ == "__main__":
    main()
