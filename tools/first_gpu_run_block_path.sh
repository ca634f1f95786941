#!/bin/bash
# First hardware run of the code written without a GPU (block SA setup, block smoothers, bench.py --config 4):
#   gpurun --timeout 1500 -- 'bash tools/first_gpu_run_block_path.sh'
# Everything lands in gpurun_out/block_path/.  Order: cheapest and most informative first, each step under its own timeout.
set -u
out=gpurun_out/block_path
mkdir -p $out
# 1. the xfail-marked tests with their real outcome (tracebacks on failure)
timeout 600 python -m pytest tests/test_gpu_zblock_setup.py -q --runxfail -x -rA > $out/zblock.log 2>&1; echo "zblock rc=$?" | tee $out/summary.txt
# 2. the rest of the GPU suite (regression check of the verified paths after the struct / library changes)
timeout 900 python -m pytest tests -m gpu -q -rxX > $out/gpu_suite.log 2>&1; echo "gpu suite rc=$?" | tee -a $out/summary.txt
tail -3 $out/gpu_suite.log | tee -a $out/summary.txt
# 3. config 4, small then full size (setup time, parity_in_run, roofline of bsr_gs_color_kernel<3>)
timeout 300 python bench.py --config 4 --grid 40 --steps 10 --warmup 3 > $out/bench_cfg4_40.json 2> $out/bench_cfg4_40.err; echo "cfg4/40 rc=$?" | tee -a $out/summary.txt
timeout 900 python bench.py --config 4 --steps 20 --warmup 3 > $out/bench_cfg4_128.json 2> $out/bench_cfg4_128.err; echo "cfg4/128 rc=$?" | tee -a $out/summary.txt
# 4. headline config unchanged?
timeout 600 python bench.py --steps 20 --warmup 3 > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $out/summary.txt
# 4b. the two-rows-in-flight row-pattern kernel (off by default): same bench with the switch on
B200AMG_ROWPAT_ILP2=1 timeout 600 python bench.py --steps 20 --warmup 3 > $out/bench_cfg2_ilp2.json 2> $out/bench_cfg2_ilp2.err; echo "cfg2 ilp2 rc=$?" | tee -a $out/summary.txt
B200AMG_ROWPAT_ILP2=1 timeout 600 python -m pytest tests/test_gpu_cycle.py tests/test_gpu_kernels.py -q -x > $out/gpu_suite_ilp2.log 2>&1; echo "gpu suite ilp2 rc=$?" | tee -a $out/summary.txt
python - <<'PY' | tee -a gpurun_out/block_path/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/block_path/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        print(f.split('/')[-1], round(d['value'], 1), d['unit'], 'e2e', round(d['e2e']['value'], 1), 'roofline', r.get('frac'), r.get('kernel', '')[:60],
              'cpu', (d.get('cpu_baseline') or {}).get('value'), 'parity', d.get('parity_in_run'), 'setup_s', d['config'].get('setup_s_device'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
# 4c. per-entry-point roofline table incl. the block smoothers
timeout 600 python tools/bench_kernels.py --out $out/kernels_roofline.json > $out/kernels_roofline.log 2>&1; echo "bench_kernels rc=$?" | tee -a $out/summary.txt
# 5. launch list of the config-4 cycle for profiles/ (only after the plain run above exited)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/ncu_launches_cfg4.csv \
    python bench.py --config 4 --steps 2 --warmup 1 --no-cpu-baseline > $out/ncu_cfg4.log 2>&1; echo "ncu rc=$?" | tee -a $out/summary.txt
