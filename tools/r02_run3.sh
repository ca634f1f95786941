#!/bin/bash
# round 2, GPU call 3: TMA-staged row-pattern sweep for real (plan from the most frequent pattern), wavefront
# Gauss-Seidel, reference-default improve_candidates (golden 256^3 history), ncu inside the timed region only
set -u
out=gpurun_out/r02c
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests -m gpu -x -q > $out/gpu_suite.log 2>&1; echo "gpu suite rc=$?" | tee -a $sum
tail -12 $out/gpu_suite.log >> $sum
sweep() {
  local label=$1; shift
  env "$@" timeout 400 python bench.py --no-cpu-baseline --steps 20 > $out/v_$label.json 2> $out/v_$label.err || { echo "$label FAILED" | tee -a $sum; tail -3 $out/v_$label.err | tee -a $sum; return; }
  python - "$label" "$out/v_$label.json" <<'PY' | tee -a $sum
import json,sys
d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
o=d["roofline"]["op_ms"]
g=d.get('parity_vs_reference_golden') or {}
print(sys.argv[1], round(d["value"],1), 'frac', d["roofline"]["frac"], {k:o[k] for k in ("presmooth@L0","residual@L0","restrict@L0","prolong@L0","postsmooth@L0")}, 'setup', d['detail']['setup_s_device'], 'golden', g.get('history_vs_reference_max_rel_diff'), g.get('levels_equal'), g.get('b_identical'), g.get('error'))
PY
}
sweep tileA X=1
sweep old B200AMG_ROWPAT_TILE=0
for v in B C D E; do sweep tile$v B200AMG_LIB=$PWD/variants/lib$v.so; done
timeout 900 python bench.py --steps 20 --warmup 3 > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $sum
timeout 900 ncu --set full --clock-control none --profile-from-start off -c 20 -o $out/cfg2_cycle \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $out/ncu_full_cfg2.log 2>&1; echo "ncu cfg2 rc=$?" | tee -a $sum
ncu -i $out/cfg2_cycle.ncu-rep --page raw --csv > $out/cfg2_cycle_raw.csv 2>/dev/null
python tools/ncu_summary.py $out/cfg2_cycle.ncu-rep > $out/cfg2_cycle_summary.md 2>/dev/null
ls -la $out/cfg2_cycle.ncu-rep | tee -a $sum
[ $(stat -c %s $out/cfg2_cycle.ncu-rep) -gt 30000000 ] && rm -f $out/cfg2_cycle.ncu-rep
timeout 900 ncu --set full --clock-control none --profile-from-start off -c 90 -o $out/cfg4_cycle \
    python bench.py --config 4 --steps 1 --warmup 1 --no-cpu-baseline > $out/ncu_full_cfg4.log 2>&1; echo "ncu cfg4 rc=$?" | tee -a $sum
ncu -i $out/cfg4_cycle.ncu-rep --page raw --csv > $out/cfg4_cycle_raw.csv 2>/dev/null
python tools/ncu_summary.py $out/cfg4_cycle.ncu-rep > $out/cfg4_cycle_summary.md 2>/dev/null
rm -f $out/cfg4_cycle.ncu-rep
timeout 1500 python bench.py --impl reference --steps 20 --warmup 5 > $out/ref_arm_256.json 2> $out/ref_arm_256.err; echo "ref arm 256 rc=$?" | tee -a $sum
du -sh gpurun_out | tee -a $sum
cat $sum
