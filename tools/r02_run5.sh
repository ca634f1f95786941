#!/bin/bash
# round 2, GPU call 5: full suite (PCG, solve_many, wavefront block sweeps, BSR warp kernels), config 4 with the
# per-level schedule rules, slice-per-CTA SELL kernel A/B on config 2
set -u
out=gpurun_out/r02e
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu suite rc=$?" | tee -a $sum
tail -30 $out/gpu_suite.log | cut -c1-300 >> $sum
for w in 1 0; do
  B200AMG_SELL_WIDE=$w timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_cfg2_wide$w.json 2> $out/bench_cfg2_wide$w.err; echo "cfg2 wide=$w rc=$?" | tee -a $sum
  tail -3 $out/bench_cfg2_wide$w.err >> $sum
done
timeout 900 python bench.py --config 4 --steps 20 --warmup 3 > $out/bench_cfg4.json 2> $out/bench_cfg4.err; echo "cfg4 rc=$?" | tee -a $sum
tail -3 $out/bench_cfg4.err >> $sum
timeout 900 python bench.py --config 3 --steps 20 --warmup 3 > $out/bench_cfg3.json 2> $out/bench_cfg3.err; echo "cfg3 rc=$?" | tee -a $sum
tail -3 $out/bench_cfg3.err >> $sum
python - <<'PY' | tee -a gpurun_out/r02e/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02e/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        print(f.split('/')[-1], round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), (d['e2e'].get('single_call') or {}).get('value'), 'frac', r.get('frac'),
              'parity', d.get('parity_in_run'), 'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'), 'cpu', (d.get('cpu_baseline') or {}).get('value'))
        print('   op_ms', r.get('op_ms'))
        print('   solve', d.get('solve'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
timeout 600 python tools/bench_kernels.py --out $out/kernels_roofline.json > $out/kernels_roofline.log 2>&1; echo "bench_kernels rc=$?" | tee -a $sum
du -sh gpurun_out | tee -a $sum
cat $sum
