#!/bin/bash
# what the driver runs at N = 2: both arms, default flags
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/bench_r3h_reference_n2.json 2> gpurun_out/bench_r3h_reference_n2.err; echo "ref rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29582 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_r3h_n2.json 2> gpurun_out/bench_r3h_n2.err; echo "n2 rc=$?"
tail -3 gpurun_out/bench_r3h_n2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_r3h_n2.json'))
print(d['value'], d['ms_per_step'], d['n_gpus'], d['scaling'], d['e2e'], d['secondary']['scenario_evals_per_sec'], d['gpu_launches'], d['state_hash'], d['clocks'])
r=json.load(open('gpurun_out/bench_r3h_reference_n2.json')); print('ref', r['value'], r['impl'], r['n_gpus'])
PY
