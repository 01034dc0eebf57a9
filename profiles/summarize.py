#!/usr/bin/env python
"""Summarises ncu outputs brought back in gpurun_out/ into the text files committed under profiles/.

    python profiles/summarize.py launches gpurun_out/launches.csv           > profiles/rNN_launches.txt
    python profiles/summarize.py full gpurun_out/prof_x.ncu-rep              > profiles/rNN_full_x.txt
"""
import collections
import csv
import io
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum', 'sm__cycles_elapsed.max',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active']


def launches(path):
    lines = [l for l in open(path) if not l.startswith('==')]
    agg = collections.OrderedDict()
    for row in csv.DictReader(io.StringIO(''.join(lines))):
        name = row['Kernel Name'].split('(')[0]
        v = float(row['Metric Value'].replace(',', ''))
        v *= {'ns': 1e-6, 'us': 1e-3, 'usecond': 1e-3, 'nsecond': 1e-6, 'ms': 1.0, 'msecond': 1.0}.get(row['Metric Unit'], 1e-6)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for k, v in agg.items() if 'dfma_peak' not in k)
    print("ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare shares)")
    print("total (excluding the fp64 peak probe) %.3f ms" % tot)
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        share = '' if 'dfma_peak' in k else "share=%5.1f%%" % (v[1] / tot * 100)
        print("%-36s launches=%4d total=%9.3f ms avg=%8.4f ms %s" % (k, v[0], v[1], v[1] / v[0], share))


def full(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    hdr, units = r[0], r[1]
    for row in r[2:]:
        print("kernel:", row[hdr.index('Kernel Name')], "id", row[0])
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print("  %-85s %s %s" % (w, row[i], units[i]))


if __name__ == '__main__':
    {'launches': launches, 'full': full}[sys.argv[1]](sys.argv[2])
