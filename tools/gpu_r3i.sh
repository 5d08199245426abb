#!/bin/bash
# conditional IF node around the re-sort kernels (LC_COND_GRAPH=1) vs plain graphs
mkdir -p gpurun_out
LC_COND_GRAPH=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_walled.py -x -q > gpurun_out/pytest_gpu_r3i_cond.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r3i_cond.log
tail -4 gpurun_out/pytest_gpu_r3i_cond.log
for c in 0 1; do
  for w in plaza1m plaza125k circle1k; do
    LC_COND_GRAPH=$c python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 --workload $w > gpurun_out/ab_r3i_c${c}_$w.json 2> gpurun_out/ab_r3i_c${c}_$w.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r3i_c${c}_$w.json')); print('cond$c', '$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], d['timeline']['entry_to_entry'])
except Exception as e: print('cond$c $w ERR', e)
PY
  done
done
tail -3 gpurun_out/ab_r3i_c1_plaza1m.err
