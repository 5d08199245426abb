#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_domain.py -x -q > gpurun_out/pytest_gpu_r3q.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r3q.log
tail -3 gpurun_out/pytest_gpu_r3q.log
for v in base tb1; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
  for w in plaza1m corridor100k; do
  python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 --workload $w > gpurun_out/ab_r3q_${v}_$w.json 2> gpurun_out/ab_r3q_${v}_$w.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r3q_${v}_$w.json')); print('$v', '$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()})
except Exception as e: print('$v ERR', e)
PY
  done
done
