#!/bin/bash
# 2-GPU box: domain tests across real peers, torchrun benches (IPC mailboxes)
mkdir -p gpurun_out
export LC_DD_TIMEOUT_MS=5000
timeout 900 python -m pytest tests/test_gpu_domain.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_dd2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dd2.log; tail -4 gpurun_out/pytest_dd2.log
python bench.py --no-cpu --no-e2e > gpurun_out/bench_plaza_n1b.json 2> gpurun_out/bench_plaza_n1b.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 20 > gpurun_out/bench_plaza_n2.json 2> gpurun_out/bench_plaza_n2.err; echo "rc=$?"; tail -3 gpurun_out/bench_plaza_n2.err
python - <<'PY'
import json
for f in ('gpurun_out/bench_plaza_n1b.json','gpurun_out/bench_plaza_n2.json'):
    try:
        d=json.load(open(f)); print(f, d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['counters'])
    except Exception as e: print(f, 'ERR', e)
PY
