#!/bin/bash
# quick-form point kernel: ECDSA GPU tests (incl. the forced second walk), then A/B against the previous build
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ecdsa.py tests/test_gpu_headline.py -m gpu -x -q 2>&1 | tail -5
bash tools/r2_run9.sh
N=41650 bash tools/r2_run9.sh
