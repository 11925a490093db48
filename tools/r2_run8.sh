#!/bin/bash
# N GPUs (N = $1): the full bench line under torchrun, as the driver's scaling run does
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L | wc -l; nproc
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench rc=$?"
grep -v "^\*\|OMP_NUM" gpurun_out/bench_${N}gpu.err | tail -5
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_${N}gpu.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'])
print('sharded',d.get('sharded_single_caller'))
print('block8mb',d['block8mb']['gpu_blocks_per_s'],'window4',d['block8mb_window']['four_lanes']['gpu_blocks_per_s'])
for k in ('chain','chain_host_table'):
    c=d.get(k)
    if c: print(k,c['gpu_blocks_per_s'],c['ms_per_block_by_phase'],c.get('host_threads'),c.get('host_memory'))
PY
