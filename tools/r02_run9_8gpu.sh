#!/bin/bash
# round 2, GPU call 9 (8 GPUs): public distributed API test (2 ranks) + the 8-GPU weak-scaling line (512^3)
set -u
out=gpurun_out/r02h
mkdir -p $out
sum=$out/summary.txt
: > $sum
nvidia-smi -L | wc -l | tee -a $sum
timeout 600 python -m pytest tests/test_gpu_distributed.py -m gpu -q -x > $out/dist_test.log 2>&1; echo "dist test rc=$?" | tee -a $sum
tail -8 $out/dist_test.log | cut -c1-300 >> $sum
for n in 8 4; do
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 20 --warmup 5 > $out/bench_${n}gpu.json 2> $out/bench_${n}gpu.err; echo "bench ${n}gpu rc=$?" | tee -a $sum
tail -4 $out/bench_${n}gpu.err | cut -c1-300 | grep -v "OMP_NUM\|\*\*\*\|Warning" >> $sum
done
python - <<'PY' | tee -a gpurun_out/r02h/summary.txt
import json
for n in (8, 4):
    try:
        d = json.loads(open('gpurun_out/r02h/bench_%dgpu.json' % n).read().strip().splitlines()[-1])
        print(n, 'gpu value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'parity', d['parity_in_run'], 'frac', d['roofline']['frac'])
        print(d['config'].get('parallelism'), d['config'].get('eager_cycle_breakdown_ms_rank0'), d['config'].get('setup_s_rank0_build_and_distribute'))
    except Exception as e:
        print(n, 'unreadable', e)
PY
cat $sum
