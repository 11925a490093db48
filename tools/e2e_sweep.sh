#!/bin/bash
# e2e chunk-size sweep for the host-pointer ECDSA path (BCP_E2E_WAVES = waves per chunk)
for w in 2 3 4 6 8; do
  BCP_E2E_WAVES=$w python bench.py --no-block --no-cpu-baseline --steps 10 --warmup 3 2>/dev/null > /tmp/e2e_$w.json
  python -c "import json; d=json.loads(open('/tmp/e2e_$w.json').read().strip().splitlines()[-1]); print('waves', $w, round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2))"
done
