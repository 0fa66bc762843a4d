#!/usr/bin/env python
"""Summarise an .ncu-rep (ncu --set full) into the small CSV kept under profiles/.
usage: ncu_summary.py <file.ncu-rep> <out.csv> ["header comment"]"""
import csv
import io
import subprocess
import sys

WANT = """gpu__time_duration.sum dram__bytes_read.sum dram__bytes_write.sum launch__registers_per_thread
launch__grid_size launch__block_size launch__waves_per_multiprocessor
sm__warps_active.avg.pct_of_peak_sustained_active sm__throughput.avg.pct_of_peak_sustained_elapsed
smsp__inst_executed.sum sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active
smsp__issue_active.avg.pct_of_peak_sustained_active gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed
sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
smsp__thread_inst_executed_per_inst_executed.ratio sm__cycles_elapsed.max smsp__cycles_active.avg
l1tex__t_bytes.sum lts__t_bytes.sum
smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio
smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio
smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio
smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio
smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio""".split()


def main():
    rep, out = sys.argv[1], sys.argv[2]
    note = sys.argv[3] if len(sys.argv) > 3 else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(out, "w") as f:
        if note:
            f.write("# " + note + "\n")
        f.write("# kernels: " + " | ".join(r[hdr.index("Kernel Name")].split("(")[0] for r in data) + "\n")
        f.write("metric,unit," + ",".join(f"launch{i}" for i in range(len(data))) + "\n")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                f.write(",".join([w, units[i]] + [r[i] for r in data]) + "\n")


if __name__ == "__main__":
    main()
