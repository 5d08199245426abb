#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L
export LC_DD_TIMEOUT_MS=3000
timeout 900 python -m pytest tests/test_gpu_domain.py -m gpu -x -q > gpurun_out/pytest_dd.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dd.log; tail -30 gpurun_out/pytest_dd.log
