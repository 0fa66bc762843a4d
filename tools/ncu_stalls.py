#!/usr/bin/env python
"""Where a kernel waits: warp-stall samples of one reason per SASS instruction and per CUDA source line, from an
.ncu-rep (ncu --set full --import-source on, -lineinfo build).
usage: ncu_stalls.py <file.ncu-rep> [reason=stall_long_sb] [top=25]"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    reason = sys.argv[2] if len(sys.argv) > 2 else "stall_long_sb"
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    hdr, line, nk, cur_file = None, None, 0, None
    per_line = collections.OrderedDict()
    per_inst = []
    total = collections.Counter()
    for r in csv.reader(io.StringIO(raw)):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Kernel Name":
            nk += 1
            if nk > 1:
                break
            continue
        if r[0] == "Line No":
            hdr = r
            ir, isamp = hdr.index(reason), hdr.index("# Samples")
            reasons = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
            continue
        if hdr is None or len(r) <= ir:
            continue
        if r[0] != "":
            line = (cur_file, r[0], r[1].strip())
            continue
        if not r[ir].isdigit():
            continue
        v = int(r[ir])
        for i, h in reasons:
            if r[i].isdigit():
                total[h] += int(r[i])
        per_line[line] = per_line.get(line, 0) + v
        per_inst.append((v, r[2], r[3].strip(), line))
    tot = sum(total.values())
    print("all stall samples by reason:", ", ".join(f"{k[6:]} {v} ({100 * v / tot:.1f}%)" for k, v in total.most_common(8)))
    print(f"--- {reason}: top source lines")
    for (f, l, s), v in sorted(per_line.items(), key=lambda kv: -kv[1])[:top]:
        if v:
            print(f"{v:7d}  {f}:{l:>4s}  {s[:110]}")
    print(f"--- {reason}: top instructions")
    for v, addr, sass, (f, l, s) in sorted(per_inst, key=lambda t: -t[0])[:top]:
        if v:
            print(f"{v:7d}  {addr}  {sass[:60]:60s} {f}:{l}")


if __name__ == "__main__":
    main()
