#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_shim.py -x -q > gpurun_out/pytest_gpu_r2i_shim.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2i_shim.log
tail -20 gpurun_out/pytest_gpu_r2i_shim.log
python bench.py --steps 20 --warmup 5 --no-cpu --no-secondary --no-e2e > gpurun_out/bench_r2i_s20.json 2> gpurun_out/bench_r2i_s20.err
python -c "
import json; d=json.load(open('gpurun_out/bench_r2i_s20.json')); print('steps20', d['value'], d['ms_per_step'], d['state_hash']['hash'])"
