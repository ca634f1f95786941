#!/bin/bash
# round 2, GPU call 12: whole smoothing steps of small levels in one launch (small_level.cu): parity + configs 3, 4
set -u
out=gpurun_out/r02k
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu suite rc=$?" | tee -a $sum
tail -30 $out/gpu_suite.log | cut -c1-300 >> $sum
for cfg in 3 4; do
  for w in 1 0; do
    B200AMG_SMALL_LEVEL=$w timeout 900 python bench.py --config $cfg --steps 20 --warmup 3 $( [ $w = 0 ] && echo --no-cpu-baseline ) > $out/bench_cfg${cfg}_small$w.json 2> $out/bench_cfg${cfg}_small$w.err; echo "cfg$cfg small=$w rc=$?" | tee -a $sum
    tail -3 $out/bench_cfg${cfg}_small$w.err >> $sum
  done
done
python - <<'PY' | tee -a gpurun_out/r02k/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02k/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'frac', r.get('frac'), 'parity', d.get('parity_in_run'))
        print('   smooth op_ms', {k: v for k, v in (r.get('op_ms') or {}).items() if 'smooth' in k})
    except Exception as e:
        print(f, 'unreadable:', e)
PY
cat $sum
