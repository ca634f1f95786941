#!/bin/bash
# diagonal-free sliced-ELL copies for the smoothers: tests and A/B on configs 2 and 3
set -u
out=gpurun_out/r02x
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1500 python -m pytest tests/test_gpu_cycle.py tests/test_gpu_kernels.py tests/test_gpu_distributed.py -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu cycle+kernels+distributed rc=$?" | tee -a $sum
tail -6 $out/gpu_suite.log | cut -c1-300 >> $sum
run() { # name, bench args, env...
  name=$1; shift
  args=$1; shift
  env "$@" timeout 900 python bench.py $args --steps 20 --warmup 3 --no-extras > $out/bench_$name.json 2> $out/bench_$name.err; echo "$name rc=$?" | tee -a $sum
}
run cfg2_default "" X=1
run cfg2_diagfree_off "--no-cpu-baseline" B200AMG_SELL_DIAGFREE=0
run cfg3_default "--config 3" X=1
run cfg3_diagfree_off "--config 3 --no-cpu-baseline" B200AMG_SELL_DIAGFREE=0
python - <<'PY' | tee -a gpurun_out/r02x/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02x/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        ops = r['op_ms']
        keep = {k: v for k, v in ops.items() if ('L0' in k or 'L1' in k or 'L2' in k) and ('smooth' in k or 'cfg2' in f)}
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'frac', r.get('frac'), 'parity', (d.get('parity_in_run') or {}).get('max_rel_diff_x_vs_cpu_reference'),
              'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'), keep)
    except Exception as e:
        print(f, 'unreadable:', e)
PY
