#!/bin/bash
# 2 GPUs: chain section only, lanes 2 and 3 (table streams at high priority); Merkle kernel tests + timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_hash.py tests/test_zz_no_alloc.py tests/test_gpu_block.py -x -q 2>&1 | tail -3
for lanes in 2 3; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$lanes bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline --only chain --chain-lanes $lanes > gpurun_out/chain2_l$lanes.json 2> gpurun_out/chain2_l$lanes.err; echo "rc=$?"; tail -2 gpurun_out/chain2_l$lanes.err
python - <<PY
import json
d=json.loads(open('gpurun_out/chain2_l$lanes.json').read().strip().splitlines()[-1])
c=d['chain']; print('lanes $lanes chain',c['gpu_blocks_per_s'],c['ms_per_block_by_phase'])
PY
done
timeout 300 python tools/secondary_probe.py > gpurun_out/plain_secondary.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_secondary.csv python tools/secondary_probe.py > gpurun_out/ncu_secondary.log 2>&1
grep -E "k_merkle|k_sha256d|k_eh_verify" gpurun_out/launches_secondary.csv | awk -F'","' '{print $5, $NF}' | tail -24
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_merkle_levels|k_sha256d_msgs" -s 10 -c 6 -o gpurun_out/prof_hash python tools/secondary_probe.py > gpurun_out/ncu_hash.log 2>&1
ls -la gpurun_out/*.ncu-rep
