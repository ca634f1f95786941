#!/bin/bash
# round 2, GPU call 4: warp-per-block-row BSR kernels (config 4 A/B), device PCG, pipelined solve_many e2e
set -u
out=gpurun_out/r02d
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests -m gpu -x -q > $out/gpu_suite.log 2>&1; echo "gpu suite rc=$?" | tee -a $sum
tail -25 $out/gpu_suite.log >> $sum
timeout 900 python bench.py --steps 20 --warmup 3 > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $sum
tail -3 $out/bench_cfg2.err >> $sum
for w in 1 0; do
  B200AMG_BSR_WARP=$w timeout 900 python bench.py --config 4 --steps 20 --warmup 3 $( [ $w = 0 ] && echo --no-cpu-baseline ) > $out/bench_cfg4_warp$w.json 2> $out/bench_cfg4_warp$w.err; echo "cfg4 warp=$w rc=$?" | tee -a $sum
  tail -3 $out/bench_cfg4_warp$w.err >> $sum
done
python - <<'PY' | tee -a gpurun_out/r02d/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02d/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        print(f.split('/')[-1], round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), (d['e2e'].get('single_call') or {}).get('value'), 'frac', r.get('frac'),
              'parity', d.get('parity_in_run'), 'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'))
        print('   op_ms', r.get('op_ms'))
        print('   solve', d.get('solve'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
timeout 600 python tools/bench_kernels.py --out $out/kernels_roofline.json > $out/kernels_roofline.log 2>&1; echo "bench_kernels rc=$?" | tee -a $sum
timeout 900 ncu --set full --clock-control none --profile-from-start off -c 120 -o $out/cfg4_cycle \
    python bench.py --config 4 --steps 1 --warmup 1 --no-cpu-baseline > $out/ncu_full_cfg4.log 2>&1; echo "ncu cfg4 rc=$?" | tee -a $sum
ncu -i $out/cfg4_cycle.ncu-rep --page raw --csv > $out/cfg4_cycle_raw.csv 2>/dev/null
python tools/ncu_summary.py $out/cfg4_cycle.ncu-rep > $out/cfg4_cycle_summary.md 2>/dev/null
rm -f $out/cfg4_cycle.ncu-rep
du -sh gpurun_out | tee -a $sum
cat $sum
