#!/bin/bash
# end-of-round check: smoke, whole GPU suite, default bench line (kernel-only legs skipped where noted)
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -12 gpurun_out/pytest_gpu.log
python bench.py --no-cpu > gpurun_out/bench_plaza.json 2> gpurun_out/bench_plaza.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_plaza.json')); print('%.3e'%d['value'], d['ms_per_step'], '%.3e'%d['e2e']['value'], d['roofline']['frac'], {k:round(v*1e3,1) for k,v in d['roofline']['kernel_ms'].items()})
PY
