#!/bin/bash
# barrier period of the v2 walk, and the persistent (rounds) form of the point kernel, at 1M signatures
mkdir -p gpurun_out
bash tools/r2_run9.sh
echo "rounds mode 1"; BCP_POINT_ROUNDS=1 timeout 300 python tools/quick_bench.py 1000000 5 2>&1 | tail -1
echo "8M"; timeout 300 python tools/quick_bench.py 8000000 3 2>&1 | tail -1
