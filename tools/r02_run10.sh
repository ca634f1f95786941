#!/bin/bash
# round 2, GPU call 10: tile kernel with a runtime ring depth (512-point grid lines), SELL slice-per-CTA kernel with the
# maximum shared-memory carve-out (A/B), strong-scaling point 512^3 on one GPU
set -u
out=gpurun_out/r02i
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu suite rc=$?" | tee -a $sum
tail -30 $out/gpu_suite.log | cut -c1-300 >> $sum
for w in 1 0; do
  B200AMG_SELL_WIDE=$w timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_cfg2_wide$w.json 2> $out/bench_cfg2_wide$w.err; echo "cfg2 wide=$w rc=$?" | tee -a $sum
  tail -3 $out/bench_cfg2_wide$w.err >> $sum
done
timeout 1500 python bench.py --grid 512 --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_512_1gpu.json 2> $out/bench_512_1gpu.err; echo "512^3 1gpu rc=$?" | tee -a $sum
tail -3 $out/bench_512_1gpu.err >> $sum
python - <<'PY' | tee -a gpurun_out/r02i/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02i/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'frac', r.get('frac'), 'setup', d['detail'].get('setup_s_device'), d['config']['levels'])
        print('   op_ms', r.get('op_ms'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
du -sh gpurun_out | tee -a $sum
cat $sum
