#!/bin/bash
# programmatic dependent launch on every kernel of the cycle: full GPU suite, then A/B on configs 2, 3, 4
set -u
out=gpurun_out/r02q
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 2400 python -m pytest tests -m gpu -q -x > $out/gpu_suite.log 2>&1; echo "full gpu suite rc=$?" | tee -a $sum
tail -5 $out/gpu_suite.log | cut -c1-300 >> $sum
run() { # name, bench args, env...
  name=$1; shift
  args=$1; shift
  env "$@" timeout 900 python bench.py $args --steps 20 --warmup 3 --no-extras > $out/bench_$name.json 2> $out/bench_$name.err; echo "$name rc=$?" | tee -a $sum
}
run cfg2_default "" X=1
run cfg2_pdl_off "--no-cpu-baseline" B200AMG_PDL=0
run cfg3_default "--config 3" X=1
run cfg3_small_level_off "--config 3 --no-cpu-baseline" B200AMG_SMALL_LEVEL=0
run cfg4_default "--config 4" X=1
run cfg4_pdl_off "--config 4 --no-cpu-baseline" B200AMG_PDL=0
python - <<'PY' | tee -a gpurun_out/r02q/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02q/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        ops = r['op_ms']
        keep = {k: v for k, v in ops.items() if ('smooth' in k and 'L0' not in k) or 'cfg2' in f}
        s = d.get('solve') or {}
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'parity', (d.get('parity_in_run') or {}).get('max_rel_diff_x_vs_cpu_reference'),
              'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'), 'e2e', (d.get('e2e') or {}).get('value'),
              'pcg', (s.get('pcg') or {}).get('ms_per_iteration'), keep)
    except Exception as e:
        print(f, 'unreadable:', e)
PY
