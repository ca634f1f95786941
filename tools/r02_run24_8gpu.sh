#!/bin/bash
# 512^3 on 8 B200 with the end-of-round kernels
set -u
out=gpurun_out/r02w
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29618 bench.py --gpus 8 --steps 20 --warmup 5 > $out/bench_8gpu.json 2> $out/bench_8gpu.err; echo "bench 8gpu rc=$?" | tee -a $sum
tail -4 $out/bench_8gpu.err | cut -c1-300 | grep -v "OMP_NUM\|\*\*\*\|Warning" >> $sum
python - <<'PY' | tee -a gpurun_out/r02w/summary.txt
import json
d = json.loads(open('gpurun_out/r02w/bench_8gpu.json').read().strip().splitlines()[-1])
print('8 gpu value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'parity', d['parity_in_run'], 'frac', d['roofline']['frac'])
print(d['config'].get('parallelism'), d['config'].get('eager_cycle_breakdown_ms_rank0'), d['config'].get('setup_s_rank0_build_and_distribute'))
PY
