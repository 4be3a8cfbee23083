"""Per-source-line instruction and stall-sample totals of one ncu capture:
   python scripts/ncu_lines.py <report.ncu-rep> [file-substring] [top]
(reads `ncu -i rep --page source --csv --print-source cuda,sass`; needs the source at the path it was compiled from)"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]; want = sys.argv[2] if len(sys.argv) > 2 else 'pcb_fast.cu'; top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
kidx = int(sys.argv[4]) if len(sys.argv) > 4 else 0      # which kernel of the report (its file sections repeat per kernel)
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur = None; hdr = None; per = collections.OrderedDict()
tot_i = tot_s = 0
first_file = None; kernel = -1
for r in rows:
    if len(r) >= 2 and r[0] in ('File Name', 'File Path'):
        cur = r[1]
        if first_file is None:
            first_file = cur
        if cur == first_file:
            kernel += 1
        continue
    if kernel != kidx:
        continue
    if len(r) > 4 and r[0] == 'Line No':
        hdr = r; continue
    if hdr is None or len(r) < len(hdr) or not r[0] or r[0] == 'Line No':
        continue
    ii = hdr.index('Instructions Executed'); si = hdr.index('# Samples')
    try:
        n_i = int(r[ii]); n_s = int(r[si])
    except ValueError:
        continue
    tot_i += n_i; tot_s += n_s
    key = (cur.split('/')[-1], int(r[0]))
    a = per.setdefault(key, [0, 0, r[1]])
    a[0] += n_i; a[1] += n_s
print('total instructions %d, samples %d' % (tot_i, tot_s))
items = sorted(per.items(), key=lambda kv: -kv[1][1])[:top]
for (f, ln), (n_i, n_s, src) in sorted(items):
    print('%-16s %5d  inst %5.2f%%  samples %5.2f%%   %s' % (f, ln, 100.0 * n_i / tot_i, 100.0 * n_s / max(1, tot_s), src.strip()[:110]))

# optional: totals per region of pcb_fast.cu given as name:lo-hi arguments after a '--'
if '--' in sys.argv:
    regs = []
    for spec in sys.argv[sys.argv.index('--') + 1:]:
        name, rng = spec.split(':'); lo, hi = rng.split('-'); regs.append((name, int(lo), int(hi)))
    acc = collections.OrderedDict((n, [0, 0]) for n, _, _ in regs); acc['other files'] = [0, 0]; acc['elsewhere'] = [0, 0]
    for (f, ln), (n_i, n_s, src) in per.items():
        key = 'other files'
        if f == 'pcb_fast.cu':
            key = 'elsewhere'
            for n, lo, hi in regs:
                if lo <= ln <= hi:
                    key = n; break
        acc[key][0] += n_i; acc[key][1] += n_s
    print('--- regions')
    for k, (n_i, n_s) in acc.items():
        print('%-14s inst %5.1f%%  samples %5.1f%%' % (k, 100.0 * n_i / tot_i, 100.0 * n_s / max(1, tot_s)))
