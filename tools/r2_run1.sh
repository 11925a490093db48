#!/bin/bash
# round-2 GPU call 1: full GPU test suite, bench line, A/B of the point-kernel launch forms and comb widths, ncu captures
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/clocks.csv &
SMI=$!
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
{
for mode in 0 1; do BCP_POINT_ROUNDS=$mode timeout 300 python tools/quick_bench.py 1000000 10; done
for mode in 0 2; do BCP_POINT_ROUNDS=$mode timeout 300 python tools/quick_bench.py 41650 20; BCP_POINT_ROUNDS=$mode timeout 300 python tools/quick_bench.py 333200 10; done
for B in 20 22; do BCP_B200_LIB=$PWD/bitcoincashplus_b200/build/variants/libbcp_b$B.so timeout 300 python tools/quick_bench.py 1000000 10; done
} > gpurun_out/ab1.log 2>&1
cat gpurun_out/ab1.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; echo "bench rc=$?"
kill $SMI
timeout 300 python tools/quick_bench.py 1000000 2 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sass__inst_executed_per_opcode --clock-control none -k regex:k_ecdsa_point -s 2 -c 1 -o gpurun_out/opc_point python tools/quick_bench.py 1000000 2 > gpurun_out/ncu_opc.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ecdsa_point -s 2 -c 1 -o gpurun_out/prof_point python tools/quick_bench.py 1000000 2 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
