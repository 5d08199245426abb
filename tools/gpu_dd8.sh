#!/bin/bash
# 8-GPU box: domain tests on GPU 0 (ranks share it) and the decomposed plaza on 8 GPUs
mkdir -p gpurun_out
export LC_DD_TIMEOUT_MS=5000
timeout 600 python -m pytest tests/test_gpu_domain.py -m gpu -x -q > gpurun_out/pytest_dd.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dd.log; tail -3 gpurun_out/pytest_dd.log
bash tools/gpu_dd_n.sh 8
