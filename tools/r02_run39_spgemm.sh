#!/bin/bash
# run 39 (the round's last GPU seconds): sort-free Galerkin rows -- A/B with bit comparison, then the GPU suite with the
# sort-free rows switched on
O=gpurun_out/r02ad; mkdir -p $O
timeout 70 python tools/spgemm_ab.py > $O/spgemm_ab.json 2> $O/spgemm_ab.err; echo "ab rc=$?" > $O/summary.txt
cat $O/spgemm_ab.json >> $O/summary.txt
B200AMG_SPGEMM_HASH=1 timeout 75 python -m pytest tests -m gpu -x -q > $O/gpu_suite_hash1.log 2>&1; echo "suite(hash=1) rc=$?" >> $O/summary.txt
tail -3 $O/gpu_suite_hash1.log >> $O/summary.txt
cat $O/summary.txt
