#!/usr/bin/env python
"""Aggregates the source page of an ncu --set full --import-source on capture by CUDA source line (needs -lineinfo).
    python profiles/by_line.py gpurun_out/prof.ncu-rep path/to/source.cuh [min_percent] [kernel name substring]
Per source line of the given file: share of the kernel's stall samples and of its executed warp instructions (the stall
reasons per kernel are in the raw page); then the opcode mix, and the samples that ncu attributes to other files (CUDA headers: barriers, shuffles)."""
import collections, csv, io, subprocess, sys
rep, srcpath = sys.argv[1], sys.argv[2]
minp = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
kfilter = sys.argv[4] if len(sys.argv) > 4 else ""
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass,cuda'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(r for r in rows if r and r[0] == 'Line No')
iL, iN, iI = hdr.index('Line No'), hdr.index('# Samples'), hdr.index('Instructions Executed')
stalls = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
by = collections.defaultdict(lambda: [0, 0, collections.Counter()])
other = collections.Counter()
ops = collections.defaultdict(lambda: [0, 0])
cur_file, cur_fn, seen_fn = srcpath, "", []
for r in rows:
    if r and r[0] == 'File Path':
        cur_file = r[1]; continue
    if r and r[0] == 'Function Name':
        cur_fn = r[1]
        if cur_fn not in seen_fn: seen_fn.append(cur_fn)
        continue
    if len(r) < len(hdr) or r[iL] == 'Line No' or kfilter not in cur_fn:
        continue
    if not r[iL].strip():
        # a SASS row: opcode statistics (each instruction appears once per file section it is attributed to; count the first)
        if cur_file.endswith(srcpath.split('/')[-1]) and len(r) > 3 and r[2].startswith('0x'):
            try:
                s_, n_ = int(r[iN] or 0), int(r[iI] or 0)
            except ValueError:
                continue
            t = r[3].split()
            op = (t[1] if t and t[0].startswith('@') and len(t) > 1 else (t[0] if t else '?')).split('.')[0]
            ops[op][0] += s_; ops[op][1] += n_
        continue
    try:
        ln, s, n = int(r[iL]), int(r[iN] or 0), int(r[iI] or 0)
    except ValueError:
        continue
    if not cur_file.endswith(srcpath.split('/')[-1]):
        other[(cur_file.split('/')[-1], ln, r[1].strip()[:60])] += s
        continue
    by[ln][0] += s; by[ln][1] += n
    for i in stalls:
        if r[i]:
            by[ln][2][hdr[i]] += int(r[i])
tot_s = sum(v[0] for v in by.values()) + sum(other.values()); tot_i = sum(v[1] for v in by.values())
src = open(srcpath).read().split('\n')
print("kernels in the capture:", "; ".join(f.split('(')[0] for f in seen_fn))
print("filter: %r   total samples %d   warp instructions %d" % (kfilter, tot_s, tot_i))
print("lines with >= %.1f %% of the samples (sample share, instruction share, source):" % minp)
for ln in sorted(by):
    s, n, st = by[ln]
    if tot_s and 100.0 * s / tot_s >= minp:
        print("%5d %5.2f%% i%5.2f%%  %s" % (ln, 100.0 * s / tot_s, 100.0 * n / max(tot_i, 1), src[ln - 1].strip()[:120] if ln <= len(src) else '?'))
print("opcode mix (sample share, instruction share):")
to_s = sum(v[0] for v in ops.values()) or 1; to_i = sum(v[1] for v in ops.values()) or 1
for op, (s, n) in sorted(ops.items(), key=lambda kv: -kv[1][1])[:18]:
    print("  %-10s samples %5.1f%%  instr %5.1f%%" % (op, 100.0 * s / to_s, 100.0 * n / to_i))
print("samples attributed to other files:")
for (fn, ln, txt), n in other.most_common(6):
    print("  %5.2f%%  %s:%d  %s" % (100.0 * n / max(tot_s, 1), fn, ln, txt))
