#!/bin/bash
# 4-GPU box: adaptive stage-2 direction count vs 64 on the decomposed plaza, plus the small workloads on one GPU
mkdir -p gpurun_out
EXTRA="--no-cpu --no-secondary --no-e2e"
for w in circle1k corridor100k plaza125k plaza1m; do
    python bench.py $EXTRA --steps 64 --warmup 16 --workload $w > gpurun_out/bench_r3a_$w.json 2> gpurun_out/bench_r3a_$w.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_r3a_$w.json')); print('$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()}, d['counters'])
except Exception as e: print('$w ERR', e)
PY
done
for k in 0 64 128; do
LC_HULL_K2=$k timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 2954$((k%10)) bench.py --gpus 4 --steps 64 --warmup 16 $EXTRA > gpurun_out/bench_r3a_n4_k$k.json 2> gpurun_out/bench_r3a_n4_k$k.err; echo "n4 k$k rc=$?"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_r3a_n4_k$k.json"))
    print('k$k', '%.3e' % d["value"], '%.1f us' % (d["ms_per_step"] * 1e3), d["state_hash"]["hash"], {k: round(v * 1e3, 1) for k, v in d["roofline"]["kernel_ms"].items()}, d['counters'], d["timeline"].get("refine_end"), d["timeline"]["ranks"][1].get("force_entry"))
except Exception as e:
    print("failed", e)
PY
done
