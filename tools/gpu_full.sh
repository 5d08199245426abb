#!/bin/bash
# What the driver does at round end on a 1-GPU box: full -m gpu suite, smoke, default bench, reference arm.
mkdir -p gpurun_out
export LC_DD_TIMEOUT_MS=5000
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_full.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_full.log; tail -6 gpurun_out/pytest_gpu_full.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py --workload pso4096 --steps 5 --warmup 3 > gpurun_out/bench_pso3.json 2> gpurun_out/bench_pso3.err; head -c 250 gpurun_out/bench_pso3.json; echo
