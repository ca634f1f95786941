#!/bin/bash
# 2 GPUs: the row-partitioned cycle's device tests (block levels new) and the 2-GPU bench line
set -u
out=gpurun_out/r02o
mkdir -p $out
sum=$out/summary.txt
: > $sum
nvidia-smi -L | tee -a $sum
timeout 900 python -m pytest tests/test_gpu_distributed.py -m gpu -q -x > $out/dist_tests.log 2>&1; echo "distributed gpu tests rc=$?" | tee -a $sum
tail -15 $out/dist_tests.log | cut -c1-400 >> $sum
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29871 bench.py --gpus 2 --steps 20 --warmup 3 > $out/bench_2gpu.json 2> $out/bench_2gpu.err; echo "bench 2gpu rc=$?" | tee -a $sum
tail -3 $out/bench_2gpu.err | cut -c1-300 >> $sum
python - <<'PY' | tee -a gpurun_out/r02o/summary.txt
import json
try:
    d = json.loads(open('gpurun_out/r02o/bench_2gpu.json').read().strip().splitlines()[-1])
    print('2gpu value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'e2e', d['e2e']['value'], 'parity', d.get('parity_in_run'))
except Exception as e:
    print('unreadable', e)
PY
