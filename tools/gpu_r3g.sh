#!/bin/bash
# 8-GPU box: stage-2 direction count sweep at N = 8, then the scale table at the driver's step counts
mkdir -p gpurun_out
EXTRA="--no-cpu --no-secondary --no-e2e"
for k in 0 64 128 256; do
LC_HULL_K2=$k timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2956$((k%7)) bench.py --gpus 8 --steps 64 --warmup 16 $EXTRA > gpurun_out/scale_r3g_s64_k${k}_n8.json 2> gpurun_out/scale_r3g_s64_k${k}_n8.err; echo "k$k rc=$?"
done
python bench.py --steps 20 --warmup 5 $EXTRA > gpurun_out/scale_r3g_n1.json 2> gpurun_out/scale_r3g_n1.err
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2957$n bench.py --gpus $n --steps 20 --warmup 5 $EXTRA > gpurun_out/scale_r3g_n$n.json 2> gpurun_out/scale_r3g_n$n.err
  echo "n=$n rc=$?"
done
python - <<PY
import json
base=None
for f in ("n1","n2","n4","n8","s64_k0_n8","s64_k64_n8","s64_k128_n8","s64_k256_n8"):
    try:
        d=json.load(open(f"gpurun_out/scale_r3g_{f}.json"))
        base=base or d["value"]
        n=d["n_gpus"]
        km={k:round(v*1e3,1) for k,v in d["roofline"]["kernel_ms"].items()}
        tl=d.get("timeline",{})
        print(f, f'{d["value"]:.3e}', f'{d["ms_per_step"]*1e3:.1f}us', 'eff', round(d["value"]/base/n,2), d["state_hash"]["hash"], km, d['counters']['candidates'], 'refine', tl.get('refine_entry'), tl.get('refine_end'), 'force_entry', (tl.get('ranks') or [{}])[0].get('force_entry'))
    except Exception as e:
        print(f,"failed",e)
PY
