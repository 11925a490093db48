#!/bin/bash
# round-2 GPU call 4 (2 GPUs): two-device chain tests, mixed blocks, script corpus, then the bench line at N=2
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_chain.py tests/test_gpu_replay.py tests/test_gpu_utxo.py -x -q > gpurun_out/pytest_2gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_2gpu.log
tail -6 gpurun_out/pytest_2gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_2gpu.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_2gpu.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'])
print('sharded',d.get('sharded_single_caller'))
for k in ('chain','chain_host_table'):
    c=d.get(k)
    if c: print(k,c['gpu_blocks_per_s'],c['ms_per_block_by_phase'])
PY
