#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2r.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2r.log
tail -5 gpurun_out/pytest_gpu_r2r.log
python bench.py > gpurun_out/bench_r2r_plaza1m.json 2> gpurun_out/bench_r2r_plaza1m.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_r2r_plaza1m.json')); print(d['value'], d['ms_per_step']); print(d['e2e']); print(d['e2e_read']); print(d['secondary']['scenario_evals_per_sec'])"
