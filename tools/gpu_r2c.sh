#!/bin/bash
# walled scenes on the GPU + the whole suite
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_walled.py -x -q > gpurun_out/pytest_gpu_r2c_walled.log 2>&1; echo "walled rc=$?" >> gpurun_out/pytest_gpu_r2c_walled.log
tail -25 gpurun_out/pytest_gpu_r2c_walled.log
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_walled.py > gpurun_out/pytest_gpu_r2c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2c.log
tail -6 gpurun_out/pytest_gpu_r2c.log
python bench.py --steps 20 --warmup 5 --no-cpu --no-secondary > gpurun_out/bench_r2c_plaza1m.json 2> gpurun_out/bench_r2c.err; python -c "
import json; d=json.load(open('gpurun_out/bench_r2c_plaza1m.json')); print(d['value'], d['ms_per_step'], d['state_hash']['hash'], d['roofline']['kernel_ms'])"
