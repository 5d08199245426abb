#!/bin/bash
# A/B of differently built library variants on the same box (LEARN_COMMS_LIB override): tools/gpu_ab.sh base x1 x2 ...
mkdir -p gpurun_out
for v in "$@"; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
  for w in plaza1m plaza125k corridor100k; do
    python bench.py --no-cpu --no-e2e --workload $w > gpurun_out/ab_${v}_$w.json 2> gpurun_out/ab_${v}_$w.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_${v}_$w.json')); print('$v', '$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()})
except Exception as e: print('$v $w ERR', e)
PY
  done
done
