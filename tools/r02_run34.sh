#!/bin/bash
# last regression of the round: full GPU suite + the config-2 line (PCG state array change, halo / distributed changes)
set -u
out=gpurun_out/r02ac
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1500 python -m pytest tests -m gpu -q > $out/gpu_suite.log 2>&1; echo "full gpu suite rc=$?" | tee -a $sum
tail -5 $out/gpu_suite.log | cut -c1-300 >> $sum
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $sum
timeout 900 python bench.py --steps 20 --warmup 3 --no-extras > $out/bench_cfg2.json 2> $out/bench_cfg2.err; echo "cfg2 rc=$?" | tee -a $sum
python - <<'PY' | tee -a gpurun_out/r02ac/summary.txt
import json
d = json.loads(open('gpurun_out/r02ac/bench_cfg2.json').read().strip().splitlines()[-1])
print(round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'frac', d['roofline']['frac'], 'golden', d['parity_vs_reference_golden']['history_vs_reference_max_rel_diff'], 'e2e', round(d['e2e']['value'], 1), 'pcg', d['solve']['pcg'])
PY
