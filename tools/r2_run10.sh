#!/bin/bash
# quick-form point kernel: ECDSA GPU tests (incl. the forced second walk), then A/B against the previous build
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ecdsa.py tests/test_gpu_headline.py -m gpu -x -q 2>&1 | tail -5
bash tools/r2_run9.sh
N=41650 bash tools/r2_run9.sh
# one profiler pass: the prep kernel of the 1M batch (the same command ran above without the profiler)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ecdsa_prep -s 1 -c 1 -o gpurun_out/prof_prep python tools/quick_bench.py 1000000 2 > gpurun_out/ncu_prep.log 2>&1
ls -la gpurun_out/prof_prep.ncu-rep
