#!/bin/bash
# PDL classes A/B (per-warp elected L1 invalidate): which early launches pay in graph replay
set -u
out=gpurun_out/r02s
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1200 python -m pytest tests/test_gpu_cycle.py tests/test_gpu_kernels.py -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu cycle+kernels rc=$?" | tee -a $sum
tail -4 $out/gpu_suite.log | cut -c1-300 >> $sum
run() { # name, bench args, env...
  name=$1; shift
  args=$1; shift
  env "$@" timeout 900 python bench.py $args --steps 20 --warmup 3 --no-extras --no-cpu-baseline > $out/bench_$name.json 2> $out/bench_$name.err; echo "$name rc=$?" | tee -a $sum
}
run cfg2_all "" B200AMG_PDL_CLASSES=7
run cfg2_sell "" B200AMG_PDL_CLASSES=2
run cfg2_other "" B200AMG_PDL_CLASSES=4
run cfg2_none "" B200AMG_PDL_CLASSES=1
run cfg3_all "--config 3" B200AMG_PDL_CLASSES=7
run cfg3_sweep "--config 3" B200AMG_PDL_CLASSES=1
run cfg3_sweep_sell "--config 3" B200AMG_PDL_CLASSES=3
run cfg4_all "--config 4" B200AMG_PDL_CLASSES=7
run cfg4_sweep "--config 4" B200AMG_PDL_CLASSES=1
run cfg4_none "--config 4" B200AMG_PDL=0
python - <<'PY' | tee -a gpurun_out/r02s/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02s/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
