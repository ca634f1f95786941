#!/bin/bash
# round 2, GPU call 7: colour-grouped BSR copy A/B on config 4, config 2 line with the extras
set -u
out=gpurun_out/r02g
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu suite rc=$?" | tee -a $sum
tail -30 $out/gpu_suite.log | cut -c1-300 >> $sum
for w in 1 0; do
  B200AMG_BSR_COLOR_COPY=$w timeout 900 python bench.py --config 4 --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_cfg4_copy$w.json 2> $out/bench_cfg4_copy$w.err; echo "cfg4 copy=$w rc=$?" | tee -a $sum
  tail -3 $out/bench_cfg4_copy$w.err >> $sum
done
timeout 1200 python bench.py --steps 20 --warmup 3 > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $sum
tail -3 $out/bench_cfg2.err >> $sum
python - <<'PY' | tee -a gpurun_out/r02g/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02g/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        print(f.split('/')[-1], round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), 'frac', r.get('frac'), 'parity', d.get('parity_in_run'))
        print('   op_ms', {k: v for k, v in (r.get('op_ms') or {}).items() if 'smooth' in k})
        if d.get('extra'):
            print('   extra', json.dumps(d['extra'])[:3000])
    except Exception as e:
        print(f, 'unreadable:', e)
PY
du -sh gpurun_out | tee -a $sum
cat $sum
