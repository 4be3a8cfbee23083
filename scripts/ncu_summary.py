#!/usr/bin/env python3
"""
Turns what a gpurun call brought back (gpurun_out/launches_<tag>.csv = ncu launch list,
gpurun_out/prof_<tag>.ncu-rep = one `ncu --set full` capture) into the tracked summaries under profiles/:

    python scripts/ncu_summary.py <tag>      ->  profiles/launches_<tag>.md, profiles/kernel_<tag>.md

Runs here (no GPU needed: `ncu -i` only reads the report).
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, 'gpurun_out')
PROF = os.path.join(ROOT, 'profiles')

RAW_KEYS = [
    'gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
    'launch__shared_mem_per_block', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
    'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
    'smsp__thread_inst_executed_per_inst_executed.ratio',
]


def launches(tag):
    path = os.path.join(OUT, 'launches_%s.csv' % tag)
    if not os.path.isfile(path):
        return None
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = [i for i, r in enumerate(rows) if r[0] == 'ID'][0]
    h = rows[hdr]
    ki, vi, ui = h.index('Kernel Name'), h.index('Metric Value'), h.index('Metric Unit')
    agg = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        v = float(r[vi].replace(',', ''))
        v = {'ns': v / 1e3, 'us': v, 'ms': v * 1e3, 's': v * 1e6}.get(r[ui], v)
        a = agg.setdefault(r[ki], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    lines = ['# ncu launch list `%s` (gpu__time_duration.sum, --clock-control none; cold-cache, serialised: shares only)' % tag, '',
             '| kernel | launches | total us | share |', '|---|---:|---:|---:|']
    for k, a in agg.items():
        lines.append('| `%s` | %d | %.1f | %.1f%% |' % (k, a[0], a[1], 100 * a[1] / tot))
    lines.append('')
    return '\n'.join(lines)


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(['ncu', '-i', rep, '--page', page, '--csv'] + list(extra), capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def kernel(tag):
    """every gpurun_out/prof_<tag>*.ncu-rep (pass 1, the two rounds launches, the scan ...) in one summary"""
    import glob
    reps = sorted(glob.glob(os.path.join(OUT, 'prof_%s*.ncu-rep' % tag)))
    if not reps:
        return None
    return '\n'.join(kernel_one(r, os.path.basename(r)[5:-8]) for r in reps)


def traffic(tag, pairs, what):
    """profiles/traffic_<tag>.json: DRAM bytes (read + write) of the captured trace kernels (pass 1 + both rounds launches) and
    the (window, adapter) pairs they processed; bench.py scales it to its own launch."""
    import glob
    import json
    total = 0.0
    found = []
    for rep in sorted(glob.glob(os.path.join(OUT, 'prof_%s_p*.ncu-rep' % tag))):
        raw = ncu_csv(rep, 'raw')
        h, units, data = raw[0], raw[1], raw[2:]
        for r in data:
            for key in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
                v, u = float(r[h.index(key)]), units[h.index(key)]
                total += v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u]
            found.append(r[h.index('Kernel Name')])
    if not found:
        return
    with open(os.path.join(PROF, 'traffic_%s.json' % tag), 'w') as f:
        json.dump({'dram_bytes': total, 'pairs': pairs, 'kernels': found, 'what': what}, f, indent=1)
    print('wrote profiles/traffic_%s.json (%.1f B per pair)' % (tag, total / pairs))


def kernel_one(rep, tag):
    raw = ncu_csv(rep, 'raw')
    h, units, data = raw[0], raw[1], raw[2:]
    ni = h.index('Kernel Name')
    lines = ['# ncu --set full capture `%s`' % tag, '']
    for r in data:
        lines += ['## `%s`' % r[ni], '', '| metric | value | unit |', '|---|---:|---|']
        for k in RAW_KEYS:
            if k in h:
                lines.append('| %s | %s | %s |' % (k, r[h.index(k)], units[h.index(k)]))
        for i, x in enumerate(h):
            if 'issue_stalled' in x and x.endswith('per_issue_active.ratio') and float(r[i] or 0) >= 0.05:
                lines.append('| %s | %s | warps/issue |' % (x.replace('smsp__average_warps_issue_stalled_', 'stall: ').replace('_per_issue_active.ratio', ''), r[i], ))
        lines.append('')
    # per-opcode executed-instruction histogram from the source page
    src = ncu_csv(rep, 'source')
    starts = [i for i, r in enumerate(src) if r and r[0] == 'Kernel Name'] + [len(src)]
    for a, b in zip(starts[:-1], starts[1:]):
        name, hh = src[a][1], src[a + 1]
        si, ei = hh.index('Source'), hh.index('Instructions Executed')
        op = collections.Counter()
        tot = 0
        for r in src[a + 2:b]:
            if len(r) > ei and r[ei].isdigit():
                m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)', r[si])
                op[m.group(2) if m else '?'] += int(r[ei])
                tot += int(r[ei])
        lines += ['## SASS opcode mix (warp instructions executed) `%s`' % name, '', 'total %d' % tot, '', '| opcode | share |', '|---|---:|']
        for k, v in op.most_common(14):
            lines.append('| %s | %.2f%% |' % (k, 100.0 * v / max(tot, 1)))
        lines.append('')
    return '\n'.join(lines)


def main():
    tag = sys.argv[1]
    os.makedirs(PROF, exist_ok=True)
    for name, text in (('launches', launches(tag)), ('kernel', kernel(tag))):
        if text:
            with open(os.path.join(PROF, '%s_%s.md' % (name, tag)), 'w') as f:
                f.write(text)
            print('wrote profiles/%s_%s.md' % (name, tag))
    if len(sys.argv) > 3:
        traffic(tag, int(sys.argv[2]), sys.argv[3])


if __name__ == '__main__':
    main()
