#!/bin/bash
# end-of-round regression: full GPU suite, smoke, the bench lines of configs 2/3/4, the reference arm, ncu launch list
# and one ncu --set full capture of the config-2 cycle
set -u
out=gpurun_out/r02v
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 2400 python -m pytest tests -m gpu -q > $out/gpu_suite.log 2>&1; echo "full gpu suite rc=$?" | tee -a $sum
tail -6 $out/gpu_suite.log | cut -c1-300 >> $sum
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $sum
tail -3 $out/smoke.log | cut -c1-300 >> $sum
timeout 1500 python bench.py --steps 20 --warmup 3 > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 (default line) rc=$?" | tee -a $sum
timeout 900 python bench.py --config 3 --steps 20 --warmup 3 > $out/bench_cfg3.json 2> $out/bench_cfg3.err; echo "cfg3 rc=$?" | tee -a $sum
timeout 900 python bench.py --config 4 --steps 20 --warmup 3 > $out/bench_cfg4.json 2> $out/bench_cfg4.err; echo "cfg4 rc=$?" | tee -a $sum
timeout 1500 python bench.py --impl reference --steps 20 --warmup 3 > $out/bench_reference_arm.json 2> $out/bench_reference_arm.err; echo "reference arm rc=$?" | tee -a $sum
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $out/launches_cfg2.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > $out/ncu_launches.log 2>&1; echo "ncu launch list rc=$?" | tee -a $sum
python tools/ncu_launches.py $out/launches_cfg2.csv > $out/launches_cfg2_summary.txt 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -c 20 -o $out/cfg2_cycle \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras > $out/ncu_full_cfg2.log 2>&1; echo "ncu full cfg2 rc=$?" | tee -a $sum
ncu -i $out/cfg2_cycle.ncu-rep --page raw --csv > $out/cfg2_cycle_raw.csv 2>/dev/null
python tools/ncu_summary.py $out/cfg2_cycle.ncu-rep > $out/cfg2_cycle_summary.md 2>/dev/null
rm -f $out/cfg2_cycle.ncu-rep
python - <<'PY' | tee -a gpurun_out/r02v/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02v/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        s = d.get('solve') or {}
        print(f.split('/')[-1], round(d['value'], 3), d['unit'], 'ms', round(d['ms_per_step'], 4), 'frac', r.get('frac'), 'parity', (d.get('parity_in_run') or {}).get('max_rel_diff_x_vs_cpu_reference'),
              'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'), 'e2e', (d.get('e2e') or {}).get('value'),
              'cpu', (d.get('cpu_baseline') or {}).get('value'), 'pcg', (s.get('pcg') or {}).get('ms_per_iteration'), 'launches', d.get('gpu_launches'), 'clocks', d.get('clocks'))
        x = d.get('extra') or {}
        if x:
            print('   extra', {k: (v.get('value') if isinstance(v, dict) else v) for k, v in x.items()})
    except Exception as e:
        print(f, 'unreadable:', e)
PY
du -sh gpurun_out | tee -a $sum
