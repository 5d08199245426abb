#!/bin/bash
# 8-GPU box: the default bench line at N = 1, 2, 4, 8 with the driver's step counts; state hashes must agree.
mkdir -p gpurun_out
TAG=${1:-r2}
EXTRA=${2:---no-cpu --no-secondary --no-e2e}
python bench.py --steps 20 --warmup 5 $EXTRA > gpurun_out/scale_${TAG}_n1.json 2> gpurun_out/scale_${TAG}_n1.err
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 $EXTRA > gpurun_out/scale_${TAG}_n$n.json 2> gpurun_out/scale_${TAG}_n$n.err
  echo "n=$n rc=$?"
done
python - <<PY
import json
base=None
for n in (1,2,4,8):
    try:
        d=json.load(open(f"gpurun_out/scale_${TAG}_n{n}.json"))
        base=base or d["value"]
        km={k:round(v*1e3,1) for k,v in d["roofline"]["kernel_ms"].items()}
        print(n, f'{d["value"]:.3e}', f'{d["ms_per_step"]*1e3:.1f}us', 'eff', round(d["value"]/base/n,2), d["state_hash"]["hash"], km)
    except Exception as e:
        print(n,"failed",e)
PY
