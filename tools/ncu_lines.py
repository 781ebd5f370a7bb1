#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` SASS dump per CUDA source line, using nvdisasm -g line markers of the
cubin the kernel came from.  usage: ncu_lines.py <sass.csv> <cubin> <mangled kernel name> [top]"""
import csv, re, subprocess, sys, collections
sass_csv, cubin, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
addr2line, cur, inside = {}, None, False
for l in txt:
    if l.startswith("//---") and ".text." in l:
        inside = (".text." + kern + " ") in l
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*);', l)
    if m:
        addr2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]
ai, ii, si, ti = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
base = None
agg = collections.defaultdict(lambda: [0, 0, 0])
for r in rows[2:]:
    if len(r) <= max(ai, ii, si):
        continue
    a = int(r[ai], 16) if r[ai].startswith("0x") else int(r[ai])
    if base is None:
        base = a
    line = addr2line.get(a - base)
    g = agg[line]
    g[0] += int(r[ii] or 0); g[1] += int(r[si] or 0); g[2] += int(r[ti] or 0)
ti_, ts_ = sum(g[0] for g in agg.values()), sum(g[1] for g in agg.values())
print("total warp-inst %.3e  samples %d" % (ti_, ts_))
src = {}
for k, g in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    text = ""
    if k:
        try:
            if k[0] not in src:
                import glob
                f = glob.glob("**/" + k[0], recursive=True)
                src[k[0]] = open(f[0]).read().splitlines() if f else []
            text = src[k[0]][k[1] - 1].strip()[:90]
        except Exception:
            pass
    print("%-22s inst %5.1f%%  samples %5.1f%%  thr/inst %4.1f  %s" % ("%s:%d" % k if k else "?", 100 * g[0] / ti_, 100 * g[1] / ts_, g[2] / max(g[0], 1), text))
