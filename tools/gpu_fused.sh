#!/bin/bash
mkdir -p gpurun_out
export LC_DD_TIMEOUT_MS=5000
timeout 900 python -m pytest tests/test_gpu_domain.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log; tail -6 gpurun_out/pytest_fused.log
python bench.py --no-cpu > gpurun_out/bench_plaza_fused.json 2> gpurun_out/bench_plaza_fused.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_plaza_fused.json'))
print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])
PY
