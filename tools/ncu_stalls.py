#!/usr/bin/env python
"""Top SASS instructions of an `ncu --page source --csv` dump by stall samples, with their stall reasons.
usage: ncu_stalls.py <sass.csv> [top]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter(); data = []
for r in rows[2:]:
    if len(r) < len(hdr): continue
    s = int(r[col["# Samples"]] or 0)
    d = {h: int(r[col[h]] or 0) for h in stalls}
    for h, v in d.items(): tot[h] += v
    data.append((s, r[col["Address"]], r[col["Source"]], int(r[col["Instructions Executed"]] or 0), d))
ts = sum(x[0] for x in data); ti = sum(x[3] for x in data)
print("samples", ts, "warp-inst %.3e" % ti)
print("stall totals:", ", ".join("%s %.1f%%" % (h[6:], 100 * v / ts) for h, v in tot.most_common() if v > ts * 0.005))
for s, a, src, n, d in sorted(data, key=lambda x: -x[0])[:top]:
    why = " ".join("%s:%d" % (h[6:], 100 * v / max(s, 1)) for h, v in sorted(d.items(), key=lambda x: -x[1])[:3] if v)
    print("%6s %5.2f%% inst %5.2f%%  %-60s %s" % (a[-5:], 100 * s / ts, 100 * n / ti, src[:60], why))
