#!/bin/bash
# 4-GPU box: domain tests on several devices + plaza-1M at N = 4 with strips balanced by local count
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_domain.py tests/test_gpu_domain_ipc.py -x -q > gpurun_out/pytest_gpu_r2x_domain_4gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2x_domain_4gpu.log
tail -3 gpurun_out/pytest_gpu_r2x_domain_4gpu.log
EXTRA="--no-cpu --no-secondary --no-e2e"
for st in 64 20; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 2953$((st%10)) bench.py --gpus 4 --steps $st --warmup $((st/4)) $EXTRA > gpurun_out/bench_r2x_n4_s$st.json 2> gpurun_out/bench_r2x_n4_s$st.err; echo "n4 s$st rc=$?"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_r2x_n4_s$st.json"))
    print('s$st', '%.3e' % d["value"], '%.1f us' % (d["ms_per_step"] * 1e3), d["state_hash"]["hash"], {k: round(v * 1e3, 1) for k, v in d["roofline"]["kernel_ms"].items()}, d["timeline"].get("nbr_wait_max"), d["timeline"].get("all_wait_max"))
except Exception as e:
    print("failed", e)
PY
done
