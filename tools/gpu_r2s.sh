#!/bin/bash
# lc_render_static: alone in a fresh process, then after every other GPU test
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_gpu_r2s_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2s_alone.log
tail -25 gpurun_out/pytest_gpu_r2s_alone.log
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2s.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2s.log
tail -6 gpurun_out/pytest_gpu_r2s.log
