#!/bin/bash
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,timestamp --format=csv,noheader,nounits | head -2
python bench.py --steps 20 --warmup 5 --no-cpu --no-secondary --no-e2e > gpurun_out/bench_r3l_n1.json 2> gpurun_out/bench_r3l_n1.err; echo "n1 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29594 bench.py --gpus 2 --steps 20 --warmup 5 --no-secondary --no-e2e > gpurun_out/bench_r3l_n2.json 2> gpurun_out/bench_r3l_n2.err; echo "n2 rc=$?"
python bench.py --workload pso4096 --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/bench_r3l_pso.json 2> gpurun_out/bench_r3l_pso.err; echo "pso rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_r3l_n1.json','gpurun_out/bench_r3l_n2.json','gpurun_out/bench_r3l_pso.json'):
    d=json.load(open(f)); print(d['n_gpus'], d['value'], d['ms_per_step'], d['clocks'])
PY
