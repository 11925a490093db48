#!/bin/bash
# larger pending-block ring: table / chain tests, then the chain section on this box's GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_utxo.py tests/test_gpu_chain.py tests/test_gpu_chain_fuzz.py -m gpu -x -q 2>&1 | tail -4
bash tools/r2_run6.sh ${1:-1}
