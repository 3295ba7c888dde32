#!/usr/bin/env python
"""Condense an Nsight Compute report into the handful of lines we track in
``profiles/``.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep "header comment" \
        > profiles/ncu_<kernel>_<version>_summary.txt

Reads the report with ``ncu -i <rep> --page raw --csv`` (works without a GPU)
and prints ``metric [unit] = value`` for the metrics below, first kernel in
the report.  The capture itself is made on the GPU box with
``ncu --set full --clock-control none --import-source on`` after the same
command has exited 0 without ncu (see ``/opt/skills/guides/B200_PROFILING.md``).
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "Kernel Name",
    "gpu__time_duration.sum",
    "sm__cycles_elapsed.avg.per_second",
    "launch__registers_per_thread",
    "launch__grid_size",
    "launch__block_size",
    "launch__shared_mem_per_block_dynamic",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
    "sm__inst_executed_pipe_uniform.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum",
    "dram__bytes_read.sum.per_second",
    "dram__bytes_write.sum",
    "lts__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
    "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
]


def main():
    if len(sys.argv) < 2:
        sys.exit(__doc__)
    rep = sys.argv[1]
    header = sys.argv[2] if len(sys.argv) > 2 else ""
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"],
                         capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    names, units = rows[0], rows[1]
    data = rows[2 + which]
    print(f"# {header} Source: {rep}")
    for want in METRICS:
        for i, n in enumerate(names):
            if n == want:
                u = f" [{units[i]}]" if units[i] else " []"
                print(f"{n}{u} = {data[i]}")
                break


if __name__ == "__main__":
    main()
