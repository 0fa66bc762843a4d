#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file ...).
usage: launch_summary.py <launches.csv> [out.txt]"""
import collections
import csv
import re
import sys


def main():
    rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    ik, iv, im = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
    iu = hdr.index("Metric Unit")
    tot, cnt = collections.Counter(), collections.Counter()
    for r in rows:
        if r is hdr or r[im] != "gpu__time_duration.sum":
            continue
        v = float(r[iv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3,
                                            "msecond": 1e3}.get(r[iu], 1.0)
        name = re.sub(r"\(.*", "", r[ik])
        tot[name] += v
        cnt[name] += 1
    all_us = sum(tot.values())
    lines = [f"{'kernel':80s} {'launches':>8s} {'total us':>12s} {'mean us':>9s} {'share':>7s}"]
    for k, v in tot.most_common():
        lines.append(f"{k[:80]:80s} {cnt[k]:8d} {v:12.1f} {v / cnt[k]:9.2f} {100 * v / all_us:6.1f}%")
    text = "\n".join(lines)
    print(text)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text + "\n")


if __name__ == "__main__":
    main()
