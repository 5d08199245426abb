#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_walled.py tests/test_ensemble.py tests/test_cli.py tests/test_pso_opt_dropin.py tests/test_gpu_batch.py -m gpu -x -q > gpurun_out/pytest_gpu_r2d.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2d.log
tail -30 gpurun_out/pytest_gpu_r2d.log
python bench.py --workload pso4096 --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_r2d_pso.json 2>gpurun_out/bench_r2d_pso.err; python -c "
import json; d=json.load(open('gpurun_out/bench_r2d_pso.json')); print(d['value'], d['scenario_evals_per_sec'], d['ms_per_step'])"
