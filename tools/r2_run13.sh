#!/bin/bash
# ncu --set full of the final point kernel at 1M signatures (one profiler pass per call; the command first runs without it)
mkdir -p gpurun_out
timeout 300 python tools/quick_bench.py 1000000 2 > gpurun_out/plain_q.log 2>&1 && tail -1 gpurun_out/plain_q.log &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ecdsa_point -s 2 -c 1 -o gpurun_out/prof_point_r2b python tools/quick_bench.py 1000000 2 > gpurun_out/ncu_full_r2b.log 2>&1
ls -la gpurun_out/prof_point_r2b.ncu-rep
