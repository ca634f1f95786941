#!/bin/bash
# block warp sweep: batches of two passes for full-wave colours (config 4 level 1), A/B
set -u
out=gpurun_out/r02ad
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests/test_gpu_zblock_setup.py tests/test_gpu_kernels.py -m gpu -q -k "bsr or block or elasticity" > $out/block_tests.log 2>&1; echo "block tests rc=$?" | tee -a $sum
tail -3 $out/block_tests.log | cut -c1-200 >> $sum
timeout 900 python bench.py --config 4 --steps 20 --warmup 3 > $out/bench_cfg4.json 2> $out/bench_cfg4.err; echo "cfg4 rc=$?" | tee -a $sum
B200AMG_BSR_WARP_BATCH=2 timeout 900 python bench.py --config 4 --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_cfg4_batch2.json 2> $out/bench_cfg4_batch2.err; echo "cfg4 (batches for sub-wave colours only) rc=$?" | tee -a $sum
python - <<'PY' | tee -a gpurun_out/r02ad/summary.txt
import json
for f in ('bench_cfg4', 'bench_cfg4_batch2'):
    d = json.loads(open('gpurun_out/r02ad/%s.json' % f).read().strip().splitlines()[-1])
    o = d['roofline']['op_ms']
    print(f, round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'parity', (d.get('parity_in_run') or {}).get('max_rel_diff_x_vs_cpu_reference'), {k: v for k, v in o.items() if 'smooth' in k})
PY
