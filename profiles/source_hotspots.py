#!/usr/bin/env python
"""Aggregates the ncu source page (SASS level with -lineinfo) by opcode class and by hottest instructions.
    python profiles/source_hotspots.py gpurun_out/prof_x.ncu-rep [kernel-id]
"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
blocks = out.split('"Kernel Name"')
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
blk = '"Kernel Name"' + blocks[1 + 2 * which]   # every kernel appears twice in the csv
lines = blk.splitlines()
r = list(csv.reader(io.StringIO('\n'.join(lines[1:]))))
hdr = r[0]
iS, iN, iI = hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot_s = tot_i = 0
by_op = collections.defaultdict(lambda: [0, 0])
rows = []
for row in r[1:]:
    if len(row) < len(hdr):
        continue
    s, n = int(row[iN] or 0), int(row[iI] or 0)
    op = row[iS].split()[0] if row[iS].split() else '?'
    if op.startswith('@'):
        op = row[iS].split()[1]
    op = op.split('.')[0]
    by_op[op][0] += s
    by_op[op][1] += n
    tot_s += s
    tot_i += n
    st = {hdr[i]: int(row[i] or 0) for i in stall_cols if int(row[i] or 0)}
    rows.append((s, n, row[iS].strip(), st))
print("total samples", tot_s, "total warp instructions", tot_i)
print("by opcode (samples%, instr%):")
for op, (s, n) in sorted(by_op.items(), key=lambda kv: -kv[1][0])[:22]:
    print("  %-10s samples %5.1f%%  instr %5.1f%%  (%d)" % (op, 100.0 * s / tot_s, 100.0 * n / tot_i, n))
print("hottest instructions:")
for s, n, src, st in sorted(rows, key=lambda t: -t[0])[:25]:
    top = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print("  %5.2f%%  x%-9d %-60s %s" % (100.0 * s / tot_s, n, src[:60], top))
