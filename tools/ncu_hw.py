#!/usr/bin/env python
"""Hardware-side counters of one kernel from an .ncu-rep -> profiles/r2_hw_counters.json[key] (read by bench.py's
roofline legs: `roofline.hw`, instructions per warp-step).
usage: ncu_hw.py <file.ncu-rep> <key> <warp_steps_per_launch> "<source note>" [launch_index=0]"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
OUT = os.path.join(ROOT, "profiles", "r2_hw_counters.json")


def main():
    rep, key, units, note = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4]
    li = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, row = rows[0], rows[2 + li]
    g = lambda name: float(row[hdr.index(name)].replace(",", ""))
    inst = g("smsp__inst_executed.sum")
    fp64_el = g("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed") / 100
    el = g("smsp__cycles_elapsed.avg")
    n_smsp = g("smsp__cycles_elapsed.sum") / el
    rec = {
        "kernel": row[hdr.index("Kernel Name")].split("(")[0],
        "duration_us": g("gpu__time_duration.sum"),
        "registers": g("launch__registers_per_thread"),
        "inst_per_warp_step": inst / units,
        # an FP64 warp instruction holds the sub-partition's FP64 pipe for two cycles (tools/ubench/fp64_latency.cu)
        "fp64_inst_per_warp_step": fp64_el * el * n_smsp / 2 / units,
        "fp64_pipe_frac_elapsed": fp64_el,
        "fp64_pipe_frac_active": g("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active") / 100,
        "issue_frac": g("sm__issue_active.avg.pct_of_peak_sustained_elapsed") / 100,
        "issue_frac_active": g("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100,
        "cycles_active_over_elapsed": g("smsp__cycles_active.avg") / el,
        "dram_bytes": g("dram__bytes_read.sum") * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}[
            rows[1][hdr.index("dram__bytes_read.sum")]] + g("dram__bytes_write.sum") * {
                "Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}[rows[1][hdr.index("dram__bytes_write.sum")]],
        "stall_long_scoreboard_per_issue": g("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
        "source": note,
    }
    data = {}
    if os.path.exists(OUT):
        with open(OUT) as fh:
            data = json.load(fh)
    data[key] = rec
    with open(OUT, "w") as fh:
        json.dump(data, fh, indent=1)
    print(key, json.dumps(rec))


if __name__ == "__main__":
    main()
