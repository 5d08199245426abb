#!/bin/bash
# 2-GPU box: domain tests across real peers, torchrun benches (IPC mailboxes)
mkdir -p gpurun_out
export LC_DD_TIMEOUT_MS=5000
timeout 900 python -m pytest tests/test_gpu_domain.py -m gpu -x -q > gpurun_out/pytest_dd2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dd2.log; tail -4 gpurun_out/pytest_dd2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 20 > gpurun_out/bench_plaza_n2.json 2> gpurun_out/bench_plaza_n2.err; echo "rc=$?"; grep -v "^\*\|OMP\|^$" gpurun_out/bench_plaza_n2.err | tail -5
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 50 --warmup 5 --replicas --no-e2e > gpurun_out/bench_plaza_n2_replicas.json 2> gpurun_out/bench_plaza_n2_replicas.err; echo "rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_plaza_n2.json','gpurun_out/bench_plaza_n2_replicas.json'):
    try:
        t=open(f).read(); d=json.loads(t[t.index('{'):]); print(f, '%.3e'%d['value'], d['ms_per_step'], d['e2e'], d['gpu_launches'], d['scaling'], t.count('\n'))
    except Exception as e: print(f, 'ERR', e)
PY
