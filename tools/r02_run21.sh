#!/bin/bash
# escape slots for 16-bit column steps (R0 of config 2), pre-wait operator loads in the block warp sweep (config 4)
set -u
out=gpurun_out/r02t
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1500 python -m pytest tests/test_gpu_cycle.py tests/test_gpu_kernels.py tests/test_gpu_zblock_setup.py -m gpu -q > $out/gpu_suite.log 2>&1; echo "gpu cycle+kernels+block rc=$?" | tee -a $sum
tail -6 $out/gpu_suite.log | cut -c1-300 >> $sum
run() { # name, bench args, env...
  name=$1; shift
  args=$1; shift
  env "$@" timeout 900 python bench.py $args --steps 20 --warmup 3 --no-extras > $out/bench_$name.json 2> $out/bench_$name.err; echo "$name rc=$?" | tee -a $sum
}
run cfg2_default "" X=1
run cfg2_col16_off "--no-cpu-baseline" B200AMG_SELL_COL16=0
run cfg4_default "--config 4" X=1
run cfg4_pdl0 "--config 4 --no-cpu-baseline" B200AMG_PDL=0
run cfg3_default "--config 3" X=1
python - <<'PY' | tee -a gpurun_out/r02t/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02t/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        ops = r['op_ms']
        keep = {k: v for k, v in ops.items() if ('cfg2' in f and ('L0' in k or 'L1' in k)) or ('cfg4' in f and 'L2' in k)}
        lay = (d.get('detail') or {}).get('engine_layouts') or []
        print(f.split('/')[-1], round(d['value'], 2), 'ms', round(d['ms_per_step'], 4), 'parity', (d.get('parity_in_run') or {}).get('max_rel_diff_x_vs_cpu_reference'),
              'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'), keep, [L.get('R') for L in lay][:2])
    except Exception as e:
        print(f, 'unreadable:', e)
PY
