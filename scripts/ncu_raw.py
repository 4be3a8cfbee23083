"""Key raw metrics of every kernel in an ncu report: python scripts/ncu_raw.py <report.ncu-rep>"""
import csv, io, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, u = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__registers_per_thread', 'launch__shared_mem_per_block', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio', 'lts__t_sector_hit_rate.pct']
for v in rows[2:]:
    print('==', v[h.index('Kernel Name')][:80])
    for i, n in enumerate(h):
        if n in want:
            print('  %-80s %s %s' % (n, v[i], u[i]))
        elif 'issue_stalled' in n and 'per_issue_active' in n and float(v[i] or 0) > 0.15:
            print('  stall %-74s %s' % (n.split('issue_stalled_')[1].split('_per_issue')[0], v[i]))
