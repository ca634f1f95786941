#!/bin/bash
set -u
out=gpurun_out/r02ae
mkdir -p $out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29876 bench.py --gpus 4 --steps 20 --warmup 3 > $out/bench_4gpu.json 2> $out/bench_4gpu.err; echo "bench 4gpu rc=$?" | tee $out/summary.txt
python - <<'PY' | tee -a gpurun_out/r02ae/summary.txt
import json
d = json.loads(open('gpurun_out/r02ae/bench_4gpu.json').read().strip().splitlines()[-1])
print('4 gpu value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'parity', d['parity_in_run'], 'frac', d['roofline']['frac'], d['config'].get('parallelism'))
PY
