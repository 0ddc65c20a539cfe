#!/usr/bin/env python
"""Opcode mix + stall samples from `ncu --page source --csv` output. usage: ncu_opmix.py src.csv [top]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
ops = collections.Counter(); samples = collections.Counter(); tot = 0; tots = 0
lines = []
for r in rows[2:]:
    if len(r) < len(hdr): continue
    src = r[ix["Source"]].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
    op = op.rstrip(";")
    n = int(r[ix["Instructions Executed"]] or 0); s = int(r[ix["# Samples"]] or 0)
    ops[op] += n; samples[op] += s; tot += n; tots += s
    lines.append((s, n, r[ix["Address"]], src, r))
print("total warp-instructions %d, samples %d" % (tot, tots))
print("%-28s %12s %7s %9s %7s" % ("opcode", "executed", "share", "samples", "share"))
for op, n in ops.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 40):
    print("%-28s %12d %6.1f%% %9d %6.1f%%" % (op, n, 100.0 * n / tot, samples[op], 100.0 * samples[op] / max(tots, 1)))
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("\nhottest instructions by samples:")
for s, n, a, src, r in sorted(lines, reverse=True)[:25]:
    st = sorted(((int(r[ix[c]] or 0), c) for c in stall_cols), reverse=True)[:2]
    print("%6d %9d  %-70s %s" % (s, n, src[:70], " ".join("%s=%d" % (c[6:], v) for v, c in st)))
