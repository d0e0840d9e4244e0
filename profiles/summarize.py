#!/usr/bin/env python3
"""Turns ncu exports into the small tracked summaries under profiles/.
  ncu -i X.ncu-rep --page raw --csv > raw.csv ;  python profiles/summarize.py full raw.csv out.csv
  python profiles/summarize.py launches launches.csv        (prints per-kernel share of the step)"""
import collections
import csv
import sys

PAT = ['Kernel Name', 'gpu__time_duration', 'dram__bytes', 'gpu__dram_throughput', 'sm__throughput', 'sm__warps_active',
       'launch__registers', 'launch__occupancy', 'launch__shared', 'launch__grid', 'launch__block', 'pipe_fma',
       'pipe_tensor', 'thread_inst_executed_per_inst', 'hit_rate', 'pcsamp_warps_issue_stalled', 'lts__throughput',
       'l1tex__throughput', 'lts__t_bytes', 'lts__t_sectors_op', 'sm__inst_executed.sum', 'sm__cycles_elapsed.max',
       'l1tex__data_bank_conflicts', 'smsp__inst_executed_pipe', 'sm__inst_executed_pipe']


def full(src, dst):
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    keep = [k for k in hdr if any(s in k for s in PAT)]
    with open(dst, 'w', newline='') as f:
        w = csv.writer(f)
        w.writerow(['metric', 'unit'] + [r[hdr.index('Kernel Name')] for r in rows[2:]])
        for k in keep:
            i = hdr.index(k)
            w.writerow([k, units[i]] + [r[i] for r in rows[2:]])


def launches(src):
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    hdr = rows[0]
    ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        v = float(r[vi].replace(',', ''))
        v *= {'ms': 1.0, 'us': 1e-3, 'ns': 1e-6, 's': 1e3}.get(r[ui], 1.0)
        agg[r[ki][:70]][0] += 1
        agg[r[ki][:70]][1] += v
    tot = sum(v[1] for v in agg.values())
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:15]:
        print('%-72s n=%4d %9.3f ms %5.1f%%' % (k, v[0], v[1], 100 * v[1] / tot))


if __name__ == '__main__':
    if sys.argv[1] == 'full':
        full(sys.argv[2], sys.argv[3])
    else:
        launches(sys.argv[2])
