#!/usr/bin/env python
"""Top stall sites of an exported ncu source page (tools/ncu_export.sh -> *.sass.csv.gz): usage sass_hot.py file [n]"""
import csv, gzip, sys
f = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(gzip.open(f, "rt") if f.endswith(".gz") else open(f)))
hdr = rows[1]
ia, isrc, ismp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = []
for r in rows[2:]:
    if len(r) != len(hdr) or not r[ia].startswith("0x"):
        if body:
            break                      # second captured launch: keep the first only
        continue
    body.append(r)
tot = sum(int(r[ismp]) for r in body)
print("instructions %d, samples %d" % (len(body), tot))
agg = {}
for r in body:
    for i, h in stall_cols:
        agg[h] = agg.get(h, 0) + int(r[i] or 0)
print("stall totals:", ", ".join("%s %.1f%%" % (h[6:], 100.0 * v / tot) for h, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > 0.01 * tot))
idx = {id(r): k for k, r in enumerate(body)}
for r in sorted(body, key=lambda r: -int(r[ismp]))[:n]:
    top = sorted(((int(r[i] or 0), h[6:]) for i, h in stall_cols), reverse=True)[:2]
    print("%5d %5.2f%% %9s  %-70s %s" % (idx[id(r)], 100.0 * int(r[ismp]) / tot, r[iex], r[isrc].strip()[:70], " ".join("%s:%d" % (h, v) for v, h in top if v)))
