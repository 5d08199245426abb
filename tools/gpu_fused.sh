#!/bin/bash
mkdir -p gpurun_out
export LC_DD_TIMEOUT_MS=5000
timeout 1200 python -m pytest tests/test_gpu_domain.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log; tail -6 gpurun_out/pytest_fused.log
python bench.py --no-cpu --no-e2e > gpurun_out/bench_plaza_defer.json 2> gpurun_out/bench_plaza_defer.err
python bench.py --no-cpu --no-e2e --workload corridor100k > gpurun_out/bench_corr_defer.json 2> gpurun_out/bench_corr_defer.err
python - <<'PY'
import json
for f in ('plaza','corr'):
    try:
        d=json.load(open(f'gpurun_out/bench_{f}_defer.json')); print(f, d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])
    except Exception as e: print(f,'ERR',e)
PY
ncu --set full --clock-control none --import-source on -k k_force -s 30 -c 1 -o gpurun_out/prof_force_defer python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_force_defer.log 2>&1
