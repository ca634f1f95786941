#!/bin/bash
set -u
out=gpurun_out/r02ab
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests/test_gpu_distributed.py -m gpu -q > $out/dist_tests.log 2>&1; echo "distributed gpu tests rc=$?" | tee -a $sum
tail -4 $out/dist_tests.log | cut -c1-300 >> $sum
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29871 bench.py --gpus 2 --steps 20 --warmup 3 > $out/bench_2gpu.json 2> $out/bench_2gpu.err; echo "bench 2gpu rc=$?" | tee -a $sum
python - <<'PY' | tee -a gpurun_out/r02ab/summary.txt
import json
d = json.loads(open('gpurun_out/r02ab/bench_2gpu.json').read().strip().splitlines()[-1])
print('2gpu value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'parity', d.get('parity_in_run'), d['config'].get('parallelism'))
PY
