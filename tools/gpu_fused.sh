#!/bin/bash
# parity (single crowd + domain decomposition) and the three crowd benches; PDL on and off
mkdir -p gpurun_out
export LC_DD_TIMEOUT_MS=5000
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_domain.py -m gpu -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log; tail -6 gpurun_out/pytest_fused.log
for pdl in 0; do
for w in plaza1m plaza125k corridor100k; do
LC_PDL=$pdl timeout 600 python bench.py --no-cpu --no-e2e --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_$w.json')); print('pdl=$pdl', '$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], {k:round(v*1e3,1) for k,v in d['roofline']['kernel_ms'].items()})
except Exception as e: print('$w ERR',e)
PY
done; done
