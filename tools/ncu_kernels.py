#!/usr/bin/env python
"""Per-kernel table of one `ncu --set full` report with many launches: for every kernel name the longest launch,
its DRAM bytes, achieved DRAM GB/s and the share of the measured HBM peak.
usage: ncu_kernels.py <report.ncu-rep> [peak_GBs]"""
import csv
import io
import json
import os
import subprocess
import sys

rep = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    peak = float(sys.argv[2]) if len(sys.argv) > 2 else float(json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    peak = 6650.0
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h, units = rows[0], rows[1]


def col(name):
    if name in h:
        return h.index(name)
    for i, c in enumerate(h):  # section-prefixed copies (e.g. FBSP.TriageCompute.<metric>)
        if c.endswith("." + name):
            return i
    return None


def scale(i, v):
    u = units[i]
    f = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
    return float(v.replace(",", "")) * f


iN, iT = col("Kernel Name"), col("gpu__time_duration.sum")
iR, iW = col("dram__bytes_read.sum"), col("dram__bytes_write.sum")
iD, iS = col("dram__throughput.avg.pct_of_peak_sustained_elapsed"), col("sm__throughput.avg.pct_of_peak_sustained_elapsed")
iG, iO = col("launch__registers_per_thread"), col("sm__warps_active.avg.pct_of_peak_sustained_active")
best, count = {}, {}
for r in rows[2:]:
    if len(r) <= iT:
        continue
    name = r[iN].split("(")[0].replace("void ", "")
    count[name] = count.get(name, 0) + 1
    t = scale(iT, r[iT])
    if name not in best or t > best[name][0]:
        g = lambda i: r[i] if i is not None else "-"
        best[name] = (t, scale(iR, r[iR]), scale(iW, r[iW]), g(iD), g(iS), g(iG), g(iO))
print("# longest launch per kernel; achieved = (dram read + write) / duration; peak = %.1f GB/s (MEASURED_PEAKS.json)" % peak)
print("%-58s %4s %9s %9s %9s %9s %7s %7s %5s %5s" % ("kernel", "n", "ms", "read MB", "write MB", "GB/s", "%peak", "ncu dram%", "sm%", "regs"))
for name, (t, rd, wr, d, s, g, o) in sorted(best.items(), key=lambda kv: -kv[1][0]):
    gbs = (rd + wr) / (t * 1e-3) / 1e9 if t > 0 else 0.0
    print("%-58s %4d %9.4f %9.1f %9.1f %9.1f %6.1f%% %8s %5s %5s" % (name[:58], count[name], t, rd / 1e6, wr / 1e6, gbs, 100 * gbs / peak, d[:5], s[:5], g))
