#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_domain.py -x -q > gpurun_out/pytest_gpu_r3f.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r3f.log
tail -3 gpurun_out/pytest_gpu_r3f.log
for k in 0 64; do
for w in plaza1m plaza125k corridor100k circle1k; do
    LC_HULL_K2=$k python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 --workload $w > gpurun_out/ab_r3f_k${k}_$w.json 2> gpurun_out/ab_r3f_k${k}_$w.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r3f_k${k}_$w.json')); print('k$k', '$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()}, d['counters']['candidates'], d['timeline'].get('refine_entry'), d['timeline'].get('refine_end'))
except Exception as e: print('k$k $w ERR', e)
PY
done
done
