#!/bin/bash
# N GPUs (N = $1): utxo/chain tests, then the chain section alone with 2 and 3 check lanes per GPU
N=${1:-4}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
nproc; lscpu | grep 'Model name'
for lanes in ${LANES:-2}; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$lanes bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline --only chain --chain-lanes $lanes > gpurun_out/chain${N}_l$lanes.json 2> gpurun_out/chain${N}_l$lanes.err; echo "rc=$?"; tail -2 gpurun_out/chain${N}_l$lanes.err
python - <<PY
import json
d=json.loads(open('gpurun_out/chain${N}_l$lanes.json').read().strip().splitlines()[-1])
c=d['chain']; print('gpus $N lanes $lanes chain',c['gpu_blocks_per_s'],c['ms_per_block_by_phase'],c.get('host_memory'))
PY
done
