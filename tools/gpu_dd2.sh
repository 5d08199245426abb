#!/bin/bash
# 2-GPU box: domain tests across real peers, torchrun benches (IPC mailboxes), pso sharding
mkdir -p gpurun_out
nvidia-smi -L; nvidia-smi topo -m | head -6
export LC_DD_TIMEOUT_MS=5000
timeout 900 python -m pytest tests/test_gpu_domain.py tests/test_gpu_batch.py -m gpu -x -q > gpurun_out/pytest_dd2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dd2.log; tail -8 gpurun_out/pytest_dd2.log
python bench.py --workload pso4096 --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_pso_n1.json 2> gpurun_out/bench_pso_n1.err; head -c 300 gpurun_out/bench_pso_n1.json; echo
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 20 > gpurun_out/bench_plaza_n2.json 2> gpurun_out/bench_plaza_n2.err; echo "rc=$?"; cat gpurun_out/bench_plaza_n2.json; tail -5 gpurun_out/bench_plaza_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload pso4096 --steps 5 --warmup 3 > gpurun_out/bench_pso_n2.json 2> gpurun_out/bench_pso_n2.err; echo "rc=$?"; head -c 400 gpurun_out/bench_pso_n2.json; tail -3 gpurun_out/bench_pso_n2.err
