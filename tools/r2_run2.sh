#!/bin/bash
# round-2 GPU call 2: full GPU suite (new: recovery / tweak-add, reference pubkey.cpp over the compat library), A/B of the
# point-kernel forms (0 = one block per 256 signatures, 1 = persistent with chunk claiming), block sizes and comb widths
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu2.log
tail -5 gpurun_out/pytest_gpu2.log
V=$PWD/bitcoincashplus_b200/build/variants
{
for rep in 1 2; do for mode in 0 1; do echo "mode $mode"; BCP_POINT_ROUNDS=$mode timeout 300 python tools/quick_bench.py 1000000 10; done; done
for mode in 0 1; do echo "mode $mode"; BCP_POINT_ROUNDS=$mode timeout 300 python tools/quick_bench.py 333200 10; BCP_POINT_ROUNDS=$mode timeout 300 python tools/quick_bench.py 8000000 3; done
echo "mode 0 tpb 128"; BCP_POINT_ROUNDS=0 BCP_POINT_TPB=128 timeout 300 python tools/quick_bench.py 1000000 10
for B in 20 22; do for mode in 0 1; do echo "bits $B mode $mode"; BCP_POINT_ROUNDS=$mode BCP_B200_LIB=$V/libbcp_b$B.so timeout 300 python tools/quick_bench.py 1000000 10; done; done
echo "mode 2 small"; BCP_POINT_ROUNDS=2 timeout 300 python tools/quick_bench.py 41650 20; BCP_POINT_ROUNDS=0 timeout 300 python tools/quick_bench.py 41650 20
} > gpurun_out/ab2.log 2>&1
cat gpurun_out/ab2.log
