#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 --no-cpu --no-secondary > gpurun_out/bench_r3k_n1.json 2> gpurun_out/bench_r3k_n1.err; echo "n1 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29593 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_r3k_n8.json 2> gpurun_out/bench_r3k_n8.err; echo "n8 rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_r3k_n1.json','gpurun_out/bench_r3k_n8.json'):
    d=json.load(open(f)); print(d['n_gpus'], d['value'], d['ms_per_step'], d['state_hash']['hash'], d['clocks'])
PY
