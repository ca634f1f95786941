#!/bin/bash
# round 2, GPU call 2: the TMA-staged row-pattern sweep (rowpat_tile.cu): parity, A/B against rowpat.cu, variants;
# ncu --set full rows for the config-2 and config-4 kernels (reports exported to CSV and deleted: gpurun_out <= 64 MiB)
set -u
out=gpurun_out/r02b
mkdir -p $out
sum=$out/summary.txt
free -g | head -2 > $sum; nproc >> $sum
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_cycle.py -m gpu -x -q > $out/gpu_kernels_cycle.log 2>&1; echo "gpu kernels+cycle rc=$?" | tee -a $sum
tail -15 $out/gpu_kernels_cycle.log >> $sum
sweep() {  # label, env assignments...
  local label=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu-baseline --steps 20 > $out/v_$label.json 2> $out/v_$label.err || { echo "$label FAILED" | tee -a $sum; tail -3 $out/v_$label.err | tee -a $sum; return; }
  python - "$label" "$out/v_$label.json" <<'PY' | tee -a $sum
import json,sys
d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
o=d["roofline"]["op_ms"]
print(sys.argv[1], round(d["value"],1), 'frac', d["roofline"]["frac"], {k:o[k] for k in ("presmooth@L0","residual@L0","restrict@L0","prolong@L0","postsmooth@L0")}, 'golden', (d.get('parity_vs_reference_golden') or {}).get('history_vs_reference_max_rel_diff'))
PY
}
sweep tileA X=1
sweep old B200AMG_ROWPAT_TILE=0
for v in B C D E; do sweep tile$v B200AMG_LIB=$PWD/variants/lib$v.so; done
# full bench line with the CPU baseline and the golden check
timeout 900 python bench.py --steps 20 --warmup 3 > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $sum
timeout 900 python bench.py --config 4 --steps 20 --warmup 3 > $out/bench_cfg4_128.json 2> $out/bench_cfg4_128.err; echo "cfg4/128 rc=$?" | tee -a $sum
timeout 600 python tools/bench_kernels.py --out $out/kernels_roofline.json > $out/kernels_roofline.log 2>&1; echo "bench_kernels rc=$?" | tee -a $sum
# reference arm on the box, small (does the installed reference import here?)
timeout 600 python bench.py --impl reference --grid 64 --steps 3 --warmup 1 > $out/ref_arm_64.json 2> $out/ref_arm_64.err; echo "ref arm 64 rc=$?" | tee -a $sum
# ncu: config 2 cycle kernels, then config 4 block kernels; CSV + summary kept, report dropped
timeout 900 ncu --set full --clock-control none -k 'regex:rowpat|sell_kernel|csr_pipe' -c 18 -o $out/cfg2_cycle \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $out/ncu_full_cfg2.log 2>&1; echo "ncu cfg2 rc=$?" | tee -a $sum
ncu -i $out/cfg2_cycle.ncu-rep --page raw --csv > $out/cfg2_cycle_raw.csv 2>/dev/null
python tools/ncu_summary.py $out/cfg2_cycle.ncu-rep > $out/cfg2_cycle_summary.md 2>/dev/null
rm -f $out/cfg2_cycle.ncu-rep
timeout 900 ncu --set full --clock-control none -k 'regex:bsr_(gs_color|matvec|jacobi)_kernel' -c 40 -o $out/cfg4_bsr \
    python bench.py --config 4 --steps 1 --warmup 1 --no-cpu-baseline > $out/ncu_full_cfg4.log 2>&1; echo "ncu cfg4 rc=$?" | tee -a $sum
ncu -i $out/cfg4_bsr.ncu-rep --page raw --csv > $out/cfg4_bsr_raw.csv 2>/dev/null
python tools/ncu_summary.py $out/cfg4_bsr.ncu-rep > $out/cfg4_bsr_summary.md 2>/dev/null
rm -f $out/cfg4_bsr.ncu-rep
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/ncu_launches_cfg4.csv \
    python bench.py --config 4 --steps 2 --warmup 1 --no-cpu-baseline > $out/ncu_cfg4.log 2>&1; echo "ncu launches cfg4 rc=$?" | tee -a $sum
du -sh gpurun_out | tee -a $sum
cat $sum
