"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) into a markdown table for profiles/.
usage: python tools/ncu_launches.py launches.csv CYCLES > profiles/rNN/ncu_launches_summary.md"""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith('=='))]
cycles = int(sys.argv[2])
h = rows[0]
ik, ig, iv = h.index('Kernel Name'), h.index('Grid Size'), h.index('Metric Value')
acc = OrderedDict()
for r in rows[1:]:
    if len(r) <= iv or r[h.index('Metric Name')] != 'gpu__time_duration.sum':
        continue
    key = (r[ik].replace('b200amg::', '').split('(')[0].replace('void ', ''), r[ig])
    t = float(r[iv].replace(',', ''))
    unit = r[h.index('Metric Unit')]
    t = t / 1000.0 if unit in ('ns', 'nsecond') else t
    acc.setdefault(key, []).append(t)
tot = sum(sum(v) for v in acc.values()) / cycles
n = sum(len(v) for v in acc.values())
print('%d launches in %d V-cycles = %d kernels per cycle; serialised cold-cache sum = %.1f us per cycle '
      '(graph replay measured by CUDA events: see the bench JSON).\n' % (n, cycles, n // cycles, tot))
print('| kernel (grid) | launches/cycle | us per launch | share of cycle |\n|---|---|---|---|')
for (k, g), v in sorted(acc.items(), key=lambda kv: -sum(kv[1])):
    print('| %s grid=%s | %d | %.1f | %.1f %% |' % (k, g, len(v) // cycles, sum(v) / len(v), 100 * sum(v) / cycles / tot))
