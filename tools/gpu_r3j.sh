#!/bin/bash
# what the driver runs at N = 8: both arms, default flags
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29591 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/bench_r3j_reference_n8.json 2> gpurun_out/bench_r3j_reference_n8.err; echo "ref rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29592 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_r3j_n8.json 2> gpurun_out/bench_r3j_n8.err; echo "n8 rc=$?"
tail -2 gpurun_out/bench_r3j_n8.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_r3j_n8.json'))
print(d['value'], d['ms_per_step'], d['n_gpus'], d['scaling'], d['e2e']['value'], d['secondary']['scenario_evals_per_sec'], d['secondary']['agent_steps_per_sec'], d['gpu_launches'], d['state_hash']['hash'], d['clocks'])
r=json.load(open('gpurun_out/bench_r3j_reference_n8.json')); print('ref', r['value'], r['impl'], r['n_gpus'])
PY
