#!/bin/bash
set -u
out=gpurun_out/r02aa
mkdir -p $out
for v in 0 1; do
  B200AMG_HALO_PDL=$v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29873 tools/halo_microbench.py 160 > $out/micro_pdl$v.log 2>&1; echo "microbench pdl=$v rc=$?" | tee -a $out/summary.txt
  grep "us per refresh" $out/micro_pdl$v.log | cut -c1-600 | tee -a $out/summary.txt
  tail -3 $out/micro_pdl$v.log | cut -c1-300 >> $out/summary.txt
done
