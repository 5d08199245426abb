#!/bin/bash
# tools/gpu_dd_n.sh N: decomposed 1M plaza on N GPUs, per-kernel times of rank 0
mkdir -p gpurun_out
N=$1
export LC_DD_TIMEOUT_MS=8000
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/dd_plaza_n$N.json 2> gpurun_out/dd_plaza_n$N.err; echo "plaza N=$N rc=$?"
python - <<PY
import json
t=open('gpurun_out/dd_plaza_n$N.json').read(); d=json.loads(t[t.index('{'):])
print(d['n_gpus'], '%.3e'%d['value'], '%.4f'%d['ms_per_step'], 'e2e', d.get('e2e'), {k:round(v*1e3,1) for k,v in d['roofline']['kernel_ms'].items()})
PY
tail -n 3 gpurun_out/dd_plaza_n$N.err | grep -v "^\*\|OMP\|^$" | tail -5
