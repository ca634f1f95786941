#!/bin/bash
set -u
out=gpurun_out/r02m
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_cycle.py -m gpu -q -x > $out/gpu_suite.log 2>&1; echo "gpu kernels+cycle rc=$?" | tee -a $sum
tail -5 $out/gpu_suite.log | cut -c1-300 >> $sum
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $sum
timeout 900 python bench.py --config 3 --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_cfg3.json 2> $out/bench_cfg3.err; echo "cfg3 rc=$?" | tee -a $sum
timeout 1500 python bench.py --grid 512 --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_512_1gpu.json 2> $out/bench_512_1gpu.err; echo "512^3 1gpu rc=$?" | tee -a $sum
python - <<'PY' | tee -a gpurun_out/r02m/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02m/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'frac', r.get('frac'), {k: v for k, v in r['op_ms'].items() if 'L0' in k or 'L7' in k or 'L8' in k})
    except Exception as e:
        print(f, 'unreadable:', e)
PY
cat $sum
