#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_batch.py tests/test_cli.py -m gpu -x -q > gpurun_out/pytest_batch.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_batch.log; tail -5 gpurun_out/pytest_batch.log
python bench.py --workload pso4096 --steps 5 --warmup 3 > gpurun_out/bench_pso2.json 2> gpurun_out/bench_pso2.err; cat gpurun_out/bench_pso2.json | head -c 700
ncu --set full --clock-control none --import-source on -k k_batch_rollout -c 1 -o gpurun_out/prof_batch2 python bench.py --workload pso4096 --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_batch2_full.log 2>&1
