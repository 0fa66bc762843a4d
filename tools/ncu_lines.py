#!/usr/bin/env python
"""Per-source-line instruction counts of the first kernel in an .ncu-rep (ncu --set full --import-source on,
built with -lineinfo).  usage: ncu_lines.py <file.ncu-rep> <warp_units> [top]   (warp_units = warps x loop trips)"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep, W = sys.argv[1], float(sys.argv[2])
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 50
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    cur_file, hdr, line, nk = None, None, None, 0
    agg = collections.OrderedDict()
    for r in csv.reader(io.StringIO(raw)):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Kernel Name":
            nk += 1
            if nk > 1:
                break
            continue
        if r[0] == "Line No":
            hdr = r
            ie = hdr.index("Instructions Executed")
            continue
        if hdr is None or len(r) <= ie:
            continue
        if r[0] != "":
            line = (cur_file, r[0], r[1].strip())
            continue
        src = r[3].strip().split()
        if not src or not r[ie].isdigit():
            continue
        op = src[1] if src[0].startswith("@") else src[0]
        a = agg.setdefault(line, [0, 0])
        a[0] += int(r[ie])
        if op.startswith(("DFMA", "DMUL", "DADD", "DSETP")):
            a[1] += int(r[ie])
    tot = sum(a[0] for a in agg.values())
    totd = sum(a[1] for a in agg.values())
    print(f"total warp-instructions {tot} ({tot / W:.1f} per unit), fp64 {totd} ({totd / W:.1f} per unit)")
    for (f, l, s), (e, d) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{f}:{l:>4s} all {e / W:6.1f} fp64 {d / W:6.1f}  {s[:100]}")


if __name__ == "__main__":
    main()
