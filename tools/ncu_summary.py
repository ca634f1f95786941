"""Summarise an ncu report (raw page CSV) into a small markdown table for profiles/.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN/summary.md"""
import csv, subprocess, sys, io

rep = sys.argv[1]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
cols = [('Kernel Name', 'kernel'), ('Grid Size', 'grid'), ('Block Size', 'block'),
        ('gpu__time_duration.sum', 'time_us'),
        ('dram__bytes_read.sum', 'dram_rd'), ('dram__bytes_write.sum', 'dram_wr'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram_%'),
        ('l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex_%'),
        ('lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l2_%'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ_%'),
        ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue_%'),
        ('launch__registers_per_thread', 'regs'),
        ('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'stall_long_sb'),
        ('smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'stall_barrier')]
idx = [(hdr.index(c), n) for c, n in cols if c in hdr]
units = rows[1]
print('| ' + ' | '.join(n + ('' if not units[i] else ' [%s]' % units[i]) for i, n in idx) + ' |')
print('|' + '---|' * len(idx))
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    out = []
    for i, n in idx:
        v = r[i]
        try:
            v = '%.4g' % float(v)
        except ValueError:
            v = v.replace('b200amg::', '')[:60]
        out.append(v)
    print('| ' + ' | '.join(out) + ' |')
