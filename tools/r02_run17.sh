#!/bin/bash
# programmatic dependent launch for the small-colour sweeps (config 3) A/B; regression lines of configs 2 and 4
set -u
out=gpurun_out/r02p
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_cycle.py tests/test_gpu_zblock_setup.py -m gpu -q -x > $out/gpu_suite.log 2>&1; echo "gpu kernels+cycle+block rc=$?" | tee -a $sum
tail -5 $out/gpu_suite.log | cut -c1-300 >> $sum
run() { # name, bench args, env...
  name=$1; shift
  args=$1; shift
  env "$@" timeout 900 python bench.py $args --steps 20 --warmup 3 --no-extras > $out/bench_$name.json 2> $out/bench_$name.err; echo "$name rc=$?" | tee -a $sum
}
run cfg3_default "--config 3" X=1
run cfg3_pdl_off "--config 3 --no-cpu-baseline" B200AMG_PDL=0
run cfg2_default "--no-cpu-baseline" X=1
run cfg4_default "--config 4 --no-cpu-baseline" X=1
python - <<'PY' | tee -a gpurun_out/r02p/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02p/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        ops = r['op_ms']
        keep = {k: v for k, v in ops.items() if ('smooth' in k and 'L0' not in k) or 'cfg2' in f}
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'parity', (d.get('parity_in_run') or {}).get('max_rel_diff_x_vs_cpu_reference'), keep)
    except Exception as e:
        print(f, 'unreadable:', e)
PY
