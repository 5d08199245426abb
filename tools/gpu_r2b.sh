#!/bin/bash
# 2-GPU box: the IPC-path domain test (one process per GPU), the in-process domain tests on two devices, and the default
# bench line at N = 1 and N = 2 with the driver's step counts (state hashes must agree).
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests/test_gpu_domain_ipc.py tests/test_gpu_domain.py -x -q > gpurun_out/pytest_gpu_r2b_domain_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2b_domain_2gpu.log
tail -8 gpurun_out/pytest_gpu_r2b_domain_2gpu.log
python bench.py --steps 20 --warmup 5 --no-cpu --no-secondary > gpurun_out/bench_r2b_n1.json 2> gpurun_out/bench_r2b_n1.err; echo "n1 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-secondary > gpurun_out/bench_r2b_n2.json 2> gpurun_out/bench_r2b_n2.err; echo "n2 rc=$?"; tail -3 gpurun_out/bench_r2b_n2.err
python - <<'PY'
import json
for n in (1, 2):
    try:
        d = json.load(open(f"gpurun_out/bench_r2b_n{n}.json"))
        print(n, d["value"], d["ms_per_step"], d["state_hash"]["hash"], d["roofline"]["kernel_ms"])
    except Exception as e:
        print(n, "failed", e)
PY
