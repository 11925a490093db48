#!/bin/bash
# round-2 final 1-GPU call: full GPU suite, bench line, launch list of the bench command
# (one profiler pass per call: the launch list of the bench command)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_final.log
tail -4 gpurun_out/pytest_gpu_final.log
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > gpurun_out/clocks_final.csv &
SMI=$!
timeout 1200 python bench.py > gpurun_out/bench_final_1gpu.json 2> gpurun_out/bench_final_1gpu.err; echo "bench rc=$?"
kill $SMI
tail -2 gpurun_out/bench_final_1gpu.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_final_reference.json 2>/dev/null; echo "ref rc=$?"
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-block > gpurun_out/plain_bench_short.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench_final.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-block > gpurun_out/ncu_launches.log 2>&1
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_final_1gpu.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'],'parity',d.get('parity_checked'))
print('block8mb',d['block8mb']['gpu_ms_per_block'],'window4l',d['block8mb_window']['four_lanes']['gpu_blocks_per_s'])
for k in ('chain','chain_host_table'):
    c=d.get(k)
    if c: print(k,c['gpu_blocks_per_s'],c['ms_per_block_by_phase'],c.get('host_memory'))
PY
