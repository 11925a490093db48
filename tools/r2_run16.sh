#!/bin/bash
# 2 GPUs: peer-access test and the chain section
mkdir -p gpurun_out
nvidia-smi topo -m 2>/dev/null | head -6
timeout 600 python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "peer or two_devices" 2>&1 | tail -3
bash tools/r2_run6.sh 2
python - <<'PY'
import json
d=json.loads(open('gpurun_out/chain2_l2.json').read().strip().splitlines()[-1])
c=d['chain']; print('gpus_in_use',c.get('gpus_in_use'),'peer',c.get('peer_access_to_table_gpu'))
PY
