#!/bin/bash
# run 40 (what is left of the round's GPU time): sort-free Galerkin rows of up to 8192 products (B200AMG_SPGEMM_HASH=3) --
# A/B with bit comparison on the config-2 hierarchy, then the GPU test files that build hierarchies / call the product, with that mode on
O=gpurun_out/r02ae; mkdir -p $O
timeout 40 python tools/spgemm_ab.py > $O/spgemm_ab.json 2> $O/spgemm_ab.err; echo "ab rc=$?" > $O/summary.txt
cat $O/spgemm_ab.json >> $O/summary.txt
B200AMG_SPGEMM_HASH=3 timeout 40 python -m pytest tests/test_gpu_cycle.py tests/test_gpu_zblock_setup.py tests/test_gpu_kernels.py \
    -m gpu -x -q > $O/gpu_suite_hash3.log 2>&1; echo "suite(hash=3) rc=$?" >> $O/summary.txt
tail -3 $O/gpu_suite_hash3.log >> $O/summary.txt
cat $O/summary.txt
