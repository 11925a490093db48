#!/bin/bash
# round-2 GPU call 3: device-resident coin table + DeviceChain tests, full suite, bench with the new chain section
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_utxo.py tests/test_gpu_chain.py -x -q > gpurun_out/pytest_utxo.log 2>&1; echo "utxo/chain pytest rc=$?" | tee -a gpurun_out/pytest_utxo.log
tail -15 gpurun_out/pytest_utxo.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu3.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu3.log
tail -5 gpurun_out/pytest_gpu3.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_b.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_b.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'])
for k in ('chain','chain_host_table'):
    c=d.get(k)
    if c: print(k,c['gpu_blocks_per_s'],c['ms_per_block_by_phase'])
PY
