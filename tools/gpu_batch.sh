#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_batch.py -m gpu -x -q > gpurun_out/pytest_batch.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_batch.log; tail -4 gpurun_out/pytest_batch.log
python bench.py --workload pso4096 --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_pso.json 2> gpurun_out/bench_pso.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_pso.json')); print('%.3e'%d['value'], d['ms_per_step'], d.get('scenario_evals_per_sec'))
PY
