#!/bin/bash
# full GPU suite + bench after: packed partner arg-max, candidate list / replan gate / id table compiled by k_collide
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2m.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2m.log
tail -5 gpurun_out/pytest_gpu_r2m.log
for w in plaza1m corridor100k circle1k; do
python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 --workload $w > gpurun_out/bench_r2m_$w.json 2> gpurun_out/bench_r2m_$w.err
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_r2m_$w.json')); print('$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()}, d['timeline']['entry_to_entry'], d['counters'])
except Exception as e: print('$w ERR', e)
PY
done
