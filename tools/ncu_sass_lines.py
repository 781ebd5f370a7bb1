#!/usr/bin/env python
"""SASS of given source lines with execution counts. usage: ncu_sass_lines.py <sass.csv> <nvdisasm -g -c dump> <kernel> line [line...]"""
import re, csv, sys
sass_csv, dump, kern = sys.argv[1:4]
want = set(int(x) for x in sys.argv[4:])
inside = False; cur = None; addr2 = []
for l in open(dump).read().splitlines():
    if l.startswith('//---') and '.text.' in l: inside = ('.text.' + kern + ' ') in l
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = int(m.group(2)) if m.group(1).endswith('neighbors.cu') else -1
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*);', l)
    if m: addr2.append((int(m.group(1), 16), cur, m.group(2)))
rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]; ai = hdr.index('Address'); ii = hdr.index('Instructions Executed'); ti = hdr.index('Thread Instructions Executed')
base = None; ex = {}
for r in rows[2:]:
    if len(r) <= ii: continue
    a = int(r[ai], 16) if r[ai].startswith('0x') else int(r[ai])
    if base is None: base = a
    ex[a - base] = (int(r[ii] or 0), int(r[ti] or 0))
tot = sum(v[0] for v in ex.values())
for a, ln, ins in addr2:
    if ln in want:
        e = ex.get(a, (0, 0)); print("%5x L%d %6.2f%% thr %4.1f  %s" % (a, ln, 100 * e[0] / tot, e[1] / max(e[0], 1), ins[:100]))
