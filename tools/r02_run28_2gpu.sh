#!/bin/bash
# 2 GPUs: halo refresh launched early (PDL, plan entries resolved before the wait), A/B
set -u
out=gpurun_out/r02z
mkdir -p $out
sum=$out/summary.txt
: > $sum
timeout 900 python -m pytest tests/test_gpu_distributed.py -m gpu -q > $out/dist_tests.log 2>&1; echo "distributed gpu tests rc=$?" | tee -a $sum
tail -4 $out/dist_tests.log | cut -c1-300 >> $sum
for v in on off; do
  if [ $v = off ]; then export B200AMG_HALO_PDL=0; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29871 bench.py --gpus 2 --steps 20 --warmup 3 > $out/bench_2gpu_$v.json 2> $out/bench_2gpu_$v.err; echo "bench 2gpu halo_pdl=$v rc=$?" | tee -a $sum
done
python - <<'PY' | tee -a gpurun_out/r02z/summary.txt
import json
for v in ('on', 'off'):
    try:
        d = json.loads(open('gpurun_out/r02z/bench_2gpu_%s.json' % v).read().strip().splitlines()[-1])
        print(v, 'value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'parity', d.get('parity_in_run'))
    except Exception as e:
        print(v, 'unreadable', e)
PY
