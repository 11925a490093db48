#!/bin/bash
# per-opcode instruction counts of the final point and prep kernels at 1M signatures (one profiler pass)
mkdir -p gpurun_out
timeout 300 python tools/quick_bench.py 1000000 2 > gpurun_out/plain_q.log 2>&1 && tail -1 gpurun_out/plain_q.log &&
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sass__inst_executed_per_opcode --clock-control none -k regex:k_ecdsa_p -s 2 -c 2 --csv --log-file gpurun_out/opcodes_r2b.csv python tools/quick_bench.py 1000000 2 > gpurun_out/ncu_opc_r2b.log 2>&1
ls -la gpurun_out/opcodes_r2b.csv
