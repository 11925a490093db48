#!/bin/bash
# A/B of point-kernel builds (libraries under oracle/_ref/libvar_*.so, built from the same sources with other -D flags):
# device-resident 1M-signature rate of each, verdicts asserted against the expected ones by tools/quick_bench.py
mkdir -p gpurun_out
N=${N:-1000000}
( unset BCP_B200_LIB; timeout 300 python tools/quick_bench.py $N 5 ) 2>&1 | tail -1
for lib in ${VARIANTS:-$(ls oracle/_ref/libvar_*.so)}; do
  BCP_B200_LIB=/root/repo/$lib timeout 300 python tools/quick_bench.py $N 5 2>&1 | tail -1
done
( unset BCP_B200_LIB; timeout 300 python tools/quick_bench.py $N 5 ) 2>&1 | tail -1
