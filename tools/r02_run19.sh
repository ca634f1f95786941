#!/bin/bash
# PDL with the L1 invalidate after the dependency wait: small launches only (default), everything (=2), off (=0)
set -u
out=gpurun_out/r02r
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1200 python -m pytest tests/test_gpu_cycle.py tests/test_gpu_kernels.py tests/test_gpu_zblock_setup.py -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu cycle+kernels+block rc=$?" | tee -a $sum
tail -8 $out/gpu_suite.log | cut -c1-300 >> $sum
B200AMG_PDL=2 timeout 1200 python -m pytest tests/test_gpu_cycle.py -m gpu -q > $out/gpu_cycle_pdl2.log 2>&1; echo "gpu cycle tests with PDL=2 rc=$?" | tee -a $sum
tail -4 $out/gpu_cycle_pdl2.log | cut -c1-300 >> $sum
run() { # name, bench args, env...
  name=$1; shift
  args=$1; shift
  env "$@" timeout 900 python bench.py $args --steps 20 --warmup 3 --no-extras > $out/bench_$name.json 2> $out/bench_$name.err; echo "$name rc=$?" | tee -a $sum
}
run cfg2_default "" X=1
run cfg2_pdl2 "--no-cpu-baseline" B200AMG_PDL=2
run cfg2_pdl0 "--no-cpu-baseline" B200AMG_PDL=0
run cfg3_default "--config 3" X=1
run cfg3_small_level_off "--config 3 --no-cpu-baseline" B200AMG_SMALL_LEVEL=0
run cfg4_default "--config 4" X=1
run cfg4_pdl0 "--config 4 --no-cpu-baseline" B200AMG_PDL=0
python - <<'PY' | tee -a gpurun_out/r02r/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02r/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        ops = r['op_ms']
        keep = {k: v for k, v in ops.items() if ('smooth' in k and 'L0' not in k and 'L1' not in k) or ('cfg2' in f and 'L0' not in k)}
        s = d.get('solve') or {}
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'parity', (d.get('parity_in_run') or {}).get('max_rel_diff_x_vs_cpu_reference'),
              'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'), 'e2e', round((d.get('e2e') or {}).get('value') or 0, 1),
              'pcg', (s.get('pcg') or {}).get('ms_per_iteration'), keep)
    except Exception as e:
        print(f, 'unreadable:', e)
PY
