#!/bin/bash
mkdir -p gpurun_out
export LC_DD_TIMEOUT_MS=4000
timeout 900 python -m pytest tests/test_gpu_domain.py -x -q > gpurun_out/pytest_gpu_r2e_domain.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2e_domain.log
tail -30 gpurun_out/pytest_gpu_r2e_domain.log
