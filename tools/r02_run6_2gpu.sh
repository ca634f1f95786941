#!/bin/bash
# round 2, GPU call 6 (2 GPUs): distributed parity test + the 2-GPU weak-scaling line
set -u
out=gpurun_out/r02f
mkdir -p $out
sum=$out/summary.txt
: > $sum
nvidia-smi -L | tee -a $sum
timeout 600 python -m pytest tests/test_gpu_distributed.py -m gpu -q -x > $out/dist_test.log 2>&1; echo "dist test rc=$?" | tee -a $sum
tail -15 $out/dist_test.log | cut -c1-400 >> $sum
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > $out/bench_2gpu.json 2> $out/bench_2gpu.err; echo "bench 2gpu rc=$?" | tee -a $sum
tail -5 $out/bench_2gpu.err | cut -c1-400 >> $sum
python - <<'PY' | tee -a gpurun_out/r02f/summary.txt
import json
try:
    d = json.loads(open('gpurun_out/r02f/bench_2gpu.json').read().strip().splitlines()[-1])
    print('2gpu value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'parity', d['parity_in_run'], 'frac', d['roofline']['frac'])
    print(d['config'].get('parallelism'), d['config'].get('eager_cycle_breakdown_ms_rank0'))
except Exception as e:
    print('unreadable', e)
PY
cat $sum
