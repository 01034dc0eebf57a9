#!/usr/bin/env python
"""Aggregates an ncu --set full --import-source on capture by CUDA source line (needs -lineinfo).
    python profiles/by_line.py gpurun_out/prof.ncu-rep path/to/source.cuh [min_percent]
Prints raw headline metrics, then per source line: share of stall samples, warp instructions per CTA-launch unit."""
import collections, csv, io, subprocess, sys
rep, srcpath = sys.argv[1], sys.argv[2]
minp = float(sys.argv[3]) if len(sys.argv) > 3 else 0.3
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(raw)))
if len(rr) >= 3:
    hdr, vals = rr[0], rr[2]
    want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
            'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
            'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__grid_size', 'smsp__inst_executed.sum',
            'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
            'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
            'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
            'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
            'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'launch__shared_mem_per_block_dynamic']
    for w in want:
        if w in hdr:
            print("  %-90s %s %s" % (w, vals[hdr.index(w)], rr[1][hdr.index(w)]))
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass,cuda'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Line No')
hdr = rows[hi]
iL, iN, iI = hdr.index('Line No'), hdr.index('# Samples'), hdr.index('Instructions Executed')
stalls = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
by = collections.defaultdict(lambda: [0, 0, collections.Counter()])
other = collections.Counter()
cur_file = srcpath
for r in rows:
    if r and r[0] == 'File Path':
        cur_file = r[1]
        continue
    if len(r) < len(hdr) or not r[iL].strip() or r[iL] == 'Line No':
        continue
    if not cur_file.endswith(srcpath.split('/')[-1]):
        # instructions attributed to another file (CUDA headers: __syncthreads, __syncwarp, shuffles, math)
        try:
            other[(cur_file.split('/')[-1], int(r[iL]), r[1].strip()[:60])] += int(r[iN] or 0)
        except ValueError:
            pass
        continue
    try:
        ln, s, n = int(r[iL]), int(r[iN] or 0), int(r[iI] or 0)
    except ValueError:
        continue
    by[ln][0] += s; by[ln][1] += n
    for i in stalls:
        if r[i]:
            by[ln][2][hdr[i]] += int(r[i])
tot_s = sum(v[0] for v in by.values()) + sum(other.values()); tot_i = sum(v[1] for v in by.values())
src = open(srcpath).read().split('\n')
print("total samples", tot_s, "warp instructions", tot_i)
for ln in sorted(by):
    s, n, st = by[ln]
    if 100.0 * s / tot_s >= minp:
        top = ", ".join("%s %d" % (k.replace('stall_', ''), v) for k, v in st.most_common(3))
        print("%5d %5.2f%% i%5.2f%%  %-100s [%s]" % (ln, 100.0 * s / tot_s, 100.0 * n / tot_i, src[ln - 1].strip()[:100] if ln <= len(src) else '?', top))
print("samples attributed to other files:")
for (fn, ln, txt), n in other.most_common(8):
    print("  %5.2f%%  %s:%d  %s" % (100.0 * n / tot_s, fn, ln, txt))
