#!/usr/bin/env python
"""Region / line breakdown of a tools/ncu_lines.py dump for neighbors.cu.  usage: ncu_regions.py <lines.txt> [min_inst_pct]"""
import re, sys
src = open('/root/repo/cornerstone_octree_b200/csrc/neighbors.cu').read().splitlines()
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
rows = []
for l in open(sys.argv[1]):
    m = re.match(r'(\S+):(\d+)\s+inst\s+([\d.]+)%\s+samples\s+([\d.]+)%\s+thr/inst\s+([\d.]+)', l)
    if m and m.group(1) == 'neighbors.cu':
        rows.append((int(m.group(2)), float(m.group(3)), float(m.group(4)), float(m.group(5))))
rows.sort()
def find(pat):
    for i, l in enumerate(src):
        if pat in l:
            return i + 1
marks = [('helpers', 1), ('prologue', find('for (uint32_t group')), ('drain', find('auto drain')), ('processLeaf', find('auto processLeaf')),
         ('stageChildren', find('auto stageChildren')), ('overlapChildren', find('auto overlapChildren')),
         ('walk', find('if (rootMask)')), ('end', len(src) + 1)]
for (n, a), (n2, b) in zip(marks, marks[1:]):
    print("%-16s inst %5.1f%% samples %5.1f%%" % (n, sum(r[1] for r in rows if a <= r[0] < b), sum(r[2] for r in rows if a <= r[0] < b)))
for ln, i, s, t in rows:
    if i >= thr or s >= 2 * thr:
        print("%4d inst %4.1f%% smp %4.1f%% thr %4.1f | %s" % (ln, i, s, t, src[ln - 1].strip()[:100]))
