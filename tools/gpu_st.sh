#!/bin/bash
# stencil sweep on the plaza workloads
mkdir -p gpurun_out
for st in 2 3 4; do for w in plaza1m plaza125k corridor100k; do
python bench.py --no-cpu --no-e2e --workload $w --stencil $st > gpurun_out/st${st}_$w.json 2> gpurun_out/st${st}_$w.err
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/st${st}_$w.json')); print('st$st', '$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], {k:round(v*1e3,1) for k,v in d['roofline']['kernel_ms'].items()})
except Exception as e: print('st$st $w ERR', e)
PY
done; done
