#!/bin/bash
# 8-GPU box: the default bench line at N = 1, 2, 4, 8 (driver's step counts, then a longer run at 4 and 8); state hashes must agree.
mkdir -p gpurun_out
TAG=${1:-r2}
EXTRA="--no-cpu --no-secondary --no-e2e"
python bench.py --steps 20 --warmup 5 $EXTRA > gpurun_out/scale_${TAG}_n1.json 2> gpurun_out/scale_${TAG}_n1.err
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 $EXTRA > gpurun_out/scale_${TAG}_n$n.json 2> gpurun_out/scale_${TAG}_n$n.err
  echo "n=$n rc=$?"
done
for n in 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 64 --warmup 16 $EXTRA > gpurun_out/scale_${TAG}_s64_n$n.json 2> gpurun_out/scale_${TAG}_s64_n$n.err
  echo "s64 n=$n rc=$?"
done
LC_REBIN_EVERY=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29538 bench.py --gpus 8 --steps 64 --warmup 16 $EXTRA > gpurun_out/scale_${TAG}_s64k1_n8.json 2> gpurun_out/scale_${TAG}_s64k1_n8.err
python - <<PY
import json
base=None
for f in ("n1","n2","n4","n8","s64_n4","s64_n8","s64k1_n8"):
    try:
        d=json.load(open(f"gpurun_out/scale_${TAG}_{f}.json"))
        base=base or d["value"]
        n=d["n_gpus"]
        km={k:round(v*1e3,1) for k,v in d["roofline"]["kernel_ms"].items()}
        tl=d.get("timeline",{})
        print(f, f'{d["value"]:.3e}', f'{d["ms_per_step"]*1e3:.1f}us', 'eff', round(d["value"]/base/n,2), d["state_hash"]["hash"], d["state_hash"]["frames"], km, 'nbr_wait_max', tl.get('nbr_wait_max'), 'all_wait_max', tl.get('all_wait_max'))
    except Exception as e:
        print(f,"failed",e)
PY
