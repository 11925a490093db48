#!/bin/bash
# per-opcode instruction counts of the final point kernel at 1M signatures (one profiler pass, report file)
mkdir -p gpurun_out
timeout 60 python tools/quick_bench.py 1000000 2 > gpurun_out/plain_q.log 2>&1 && tail -1 gpurun_out/plain_q.log &&
timeout 60 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sass__inst_executed_per_opcode --clock-control none -k regex:k_ecdsa_point -s 2 -c 1 -o gpurun_out/opc_point_r2b python tools/quick_bench.py 1000000 2 > gpurun_out/ncu_opc_r2b.log 2>&1
ls -la gpurun_out/opc_point_r2b.ncu-rep
