#!/bin/bash
# escape-aware loop only for slices that hold escapes
set -u
out=gpurun_out/r02u
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1500 python -m pytest tests/test_gpu_cycle.py tests/test_gpu_kernels.py -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu cycle+kernels rc=$?" | tee -a $sum
tail -4 $out/gpu_suite.log | cut -c1-300 >> $sum
run() { # name, bench args, env...
  name=$1; shift
  args=$1; shift
  env "$@" timeout 900 python bench.py $args --steps 20 --warmup 3 --no-extras --no-cpu-baseline > $out/bench_$name.json 2> $out/bench_$name.err; echo "$name rc=$?" | tee -a $sum
}
run cfg2_default "" X=1
run cfg2_again "" X=2
run cfg3_default "--config 3" X=1
python - <<'PY' | tee -a gpurun_out/r02u/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02u/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        ops = r['op_ms']
        keep = {k: v for k, v in ops.items() if ('L0' in k or 'L1' in k)}
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'), keep)
    except Exception as e:
        print(f, 'unreadable:', e)
PY
