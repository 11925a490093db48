#!/bin/bash
# division-step inversion in the prep kernel: ECDSA GPU tests, then signatures-per-prep-thread sweep at 1M, 333k and one block's 41,650
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ecdsa.py tests/test_gpu_headline.py -m gpu -x -q 2>&1 | tail -4
for n in 1000000 333200 41650; do
  for k in 2 4 8 16 32; do
    echo -n "K=$k "; BCP_PREP_K=$k timeout 300 python tools/quick_bench.py $n 5 2>&1 | tail -1
  done
done
