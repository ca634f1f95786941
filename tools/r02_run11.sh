#!/bin/bash
set -u
out=gpurun_out/r02j
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_cycle.py -m gpu -q -x > $out/gpu_suite.log 2>&1; echo "gpu kernels+cycle rc=$?" | tee -a $sum
tail -5 $out/gpu_suite.log | cut -c1-300 >> $sum
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $sum
python - <<'PY' | tee -a gpurun_out/r02j/summary.txt
import json
d = json.loads(open('gpurun_out/r02j/bench_cfg2.json').read().strip().splitlines()[-1])
print(round(d['value'], 1), 'frac', d['roofline']['frac'], {k: v for k, v in d['roofline']['op_ms'].items() if 'L0' in k})
PY
cat $sum
