#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_walled.py tests/test_gpu_domain.py tests/test_gpu_batch.py -x -q > gpurun_out/pytest_gpu_r2z.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2z.log
tail -4 gpurun_out/pytest_gpu_r2z.log
for w in plaza1m corridor100k circle1k plaza125k; do
    python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 --workload $w > gpurun_out/bench_r2z_$w.json 2> gpurun_out/bench_r2z_$w.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_r2z_$w.json')); print('$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()}, d['counters'], d['timeline'].get('refine_end'))
except Exception as e: print('$w ERR', e)
PY
done
