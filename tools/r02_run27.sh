#!/bin/bash
# device PCG: one CUDA graph per iteration (A/B)
set -u
out=gpurun_out/r02y
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1200 python -m pytest tests/test_gpu_cycle.py -m gpu -q > $out/gpu_cycle.log 2>&1; echo "gpu cycle tests rc=$?" | tee -a $sum
tail -6 $out/gpu_cycle.log | cut -c1-300 >> $sum
B200AMG_PCG_GRAPH=0 timeout 600 python -m pytest tests/test_gpu_cycle.py -m gpu -q -k "pcg or preconditioner" > $out/gpu_cycle_nograph.log 2>&1; echo "pcg tests without iteration graphs rc=$?" | tee -a $sum
tail -3 $out/gpu_cycle_nograph.log | cut -c1-300 >> $sum
run() { name=$1; shift; args=$1; shift
  env "$@" timeout 900 python bench.py $args --steps 20 --warmup 3 --no-extras --no-cpu-baseline > $out/bench_$name.json 2> $out/bench_$name.err; echo "$name rc=$?" | tee -a $sum; tail -2 $out/bench_$name.err | cut -c1-300 >> $sum; }
run cfg2_default "" X=1
run cfg2_pcg_nograph "" B200AMG_PCG_GRAPH=0
run cfg3_default "--config 3" X=1
python - <<'PY' | tee -a gpurun_out/r02y/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02y/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'solve', d.get('solve'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
