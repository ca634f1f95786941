#!/bin/bash
# A/B: loads of a round forced in flight (few-row sliced-ELL forms, colour sweeps, block warp kernels), 16-bit column
# steps, small-level step with next-colour prefetch
set -u
out=gpurun_out/r02n
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
run cfg2_default "--no-cpu-baseline" X=1
run cfg2_small0_round1 "--no-cpu-baseline" B200AMG_SELL_SMALL=0
run cfg2_small1_u8_tied "--no-cpu-baseline" B200AMG_SELL_SMALL=1
run cfg2_small3_u32_tied "--no-cpu-baseline" B200AMG_SELL_SMALL=3
run cfg2_small2_tpb256 "--no-cpu-baseline" B200AMG_SELL_SMALL_TPB=256
run cfg2_col16_off "--no-cpu-baseline" B200AMG_SELL_COL16=0
run cfg3_default "--config 3" X=1
run cfg3_gs_u8 "--config 3 --no-cpu-baseline" B200AMG_SELL_GS_U=8
run cfg3_col16_off "--config 3 --no-cpu-baseline" B200AMG_SELL_COL16=0
run cfg4_default "--config 4" X=1
run cfg4_batch_off "--config 4 --no-cpu-baseline" B200AMG_BSR_WARP_BATCH=0
python - <<'PY' | tee -a gpurun_out/r02n/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02n/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        g = d.get('parity_vs_reference_golden') or {}
        lay = (d.get('detail') or {}).get('engine_layouts')
        ops = r['op_ms']
        if 'cfg2' in f:
            ops = {k: v for k, v in ops.items() if 'L0' not in k or 'restrict' in k}
        print(f.split('/')[-1], round(d['value'], 2), 'parity', (d.get('parity_in_run') or {}).get('max_rel_diff_x_vs_cpu_reference'), 'ms', round(d['ms_per_step'], 4), 'frac', r.get('frac'), 'golden', g.get('history_vs_reference_max_rel_diff'), ops)
        if 'default' in f:
            print('   layouts', lay)
    except Exception as e:
        print(f, 'unreadable:', e)
PY
cat $sum
