#!/bin/bash
# round 2, GPU call 1: unmasked block-path tests, the allocation-free entry point, config 4 on hardware, ILP2 A/B,
# ncu rows for the BSR kernels.  gpurun --timeout 1700 -- 'bash tools/r02_run1.sh'
set -u
out=gpurun_out/r02a
mkdir -p $out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $out/noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee $out/summary.txt
timeout 900 python -m pytest tests -m gpu -x -q > $out/gpu_suite.log 2>&1; echo "gpu suite rc=$?" | tee -a $out/summary.txt
tail -3 $out/gpu_suite.log | tee -a $out/summary.txt
timeout 300 python bench.py --config 4 --grid 40 --steps 10 --warmup 3 > $out/bench_cfg4_40.json 2> $out/bench_cfg4_40.err; echo "cfg4/40 rc=$?" | tee -a $out/summary.txt
timeout 900 python bench.py --config 4 --steps 20 --warmup 3 > $out/bench_cfg4_128.json 2> $out/bench_cfg4_128.err; echo "cfg4/128 rc=$?" | tee -a $out/summary.txt
timeout 600 python bench.py --steps 20 --warmup 3 > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $out/summary.txt
B200AMG_ROWPAT_ILP2=1 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_cfg2_ilp2.json 2> $out/bench_cfg2_ilp2.err; echo "cfg2 ilp2 rc=$?" | tee -a $out/summary.txt
B200AMG_ROWPAT_ILP2=1 timeout 600 python -m pytest tests/test_gpu_cycle.py tests/test_gpu_kernels.py -m gpu -q -x > $out/gpu_suite_ilp2.log 2>&1; echo "gpu suite ilp2 rc=$?" | tee -a $out/summary.txt
python - <<'PY' | tee -a gpurun_out/r02a/summary.txt
import json, glob
for f in sorted(glob.glob('gpurun_out/r02a/bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get('roofline') or {}
        print(f.split('/')[-1], round(d['value'], 1), d['unit'], 'e2e', round(d['e2e']['value'], 1), 'roofline', r.get('frac'), r.get('kernel', '')[:60],
              'cpu', (d.get('cpu_baseline') or {}).get('value'), 'parity', d.get('parity_in_run'), 'setup_s', d['config'].get('setup_s_device'))
        print('   op_ms', r.get('op_ms'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
timeout 600 python tools/bench_kernels.py --out $out/kernels_roofline.json > $out/kernels_roofline.log 2>&1; echo "bench_kernels rc=$?" | tee -a $out/summary.txt
# launch list of the config-4 cycle, then one --set full pass over the block kernels of that cycle
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/ncu_launches_cfg4.csv \
    python bench.py --config 4 --steps 2 --warmup 1 --no-cpu-baseline > $out/ncu_cfg4.log 2>&1; echo "ncu launches rc=$?" | tee -a $out/summary.txt
timeout 900 ncu --set full --clock-control none -k 'regex:bsr_(gs_color|matvec|jacobi)_kernel' -c 48 -o $out/cfg4_bsr \
    python bench.py --config 4 --steps 1 --warmup 1 --no-cpu-baseline > $out/ncu_full_cfg4.log 2>&1; echo "ncu full rc=$?" | tee -a $out/summary.txt
ncu -i $out/cfg4_bsr.ncu-rep --page raw --csv > $out/cfg4_bsr_raw.csv 2>/dev/null
python tools/ncu_summary.py $out/cfg4_bsr.ncu-rep > $out/cfg4_bsr_summary.md 2>/dev/null
ls -la $out | tee -a $out/summary.txt
