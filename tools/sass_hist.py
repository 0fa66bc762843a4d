#!/usr/bin/env python
"""Opcode histogram of address ranges of a SASS listing (one instruction per line, '/*addr*/ OP ...').
usage: sass_hist.py <listing> <lo:hi> [<lo:hi> ...]   (hex addresses, inclusive) -- the ranges of the hot path"""
import collections
import re
import sys

L = [l.rstrip() for l in open(sys.argv[1])]
idx = {}
for i, l in enumerate(L):
    m = re.match(r"/\*([0-9a-f]+)\*/", l)
    if m:
        idx[int(m.group(1), 16)] = i
cnt, tot = collections.Counter(), 0
for r in sys.argv[2:]:
    a, b = (int(x, 16) for x in r.split(":"))
    for l in L[idx[a]:idx[b] + 1]:
        m = re.match(r"/\*[0-9a-f]+\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", l)
        op = m.group(2).split(".")[0]
        if op == "IMAD" and "MOV" in m.group(2):
            op = "IMAD.MOV"
        cnt[op] += 1
        tot += 1
fp64 = sum(v for k, v in cnt.items() if k in ("DFMA", "DMUL", "DADD", "DSETP"))
print("instructions", tot, "fp64", fp64, "issue slots (fp64 x2)", tot + fp64)
print("  ".join(f"{k} {v}" for k, v in cnt.most_common()))
