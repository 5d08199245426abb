#!/bin/bash
# N-GPU box: strong scaling of the decomposed 1M plaza and of the PSO batch at N = 1, 2, 4, 8 (whatever the box has)
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
export LC_DD_TIMEOUT_MS=8000
for N in 1 2 4 8; do
  [ $N -gt $NG ] && continue
  if [ $N -eq 1 ]; then
    python bench.py --no-cpu > gpurun_out/scale_plaza_n1.json 2> gpurun_out/scale_plaza_n1.err
    python bench.py --workload pso4096 --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_pso_n1.json 2> gpurun_out/scale_pso_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/scale_plaza_n$N.json 2> gpurun_out/scale_plaza_n$N.err; echo "plaza N=$N rc=$?"
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+N)) bench.py --gpus $N --workload pso4096 --steps 5 --warmup 3 > gpurun_out/scale_pso_n$N.json 2> gpurun_out/scale_pso_n$N.err; echo "pso N=$N rc=$?"
  fi
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/scale_*.json')):
    try:
        t=open(f).read(); d=json.loads(t[t.index('{'):])
        print(f, d['n_gpus'], '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d.get('scenario_evals_per_sec'), d['roofline'].get('kernel_ms'))
    except Exception as e: print(f,'ERR',e)
PY
tail -q -n 3 gpurun_out/scale_plaza_n*.err | grep -v "^\*\|OMP\|^$" | tail -12
