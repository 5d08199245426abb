#!/bin/bash
mkdir -p gpurun_out
for v in base k128 k256; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
  for w in plaza1m corridor100k; do
    python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 --workload $w > gpurun_out/ab_r2y_${v}_$w.json 2> gpurun_out/ab_r2y_${v}_$w.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r2y_${v}_$w.json')); print('$v', '$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()}, d['counters'], d['timeline'].get('refine_end'))
except Exception as e: print('$v $w ERR', e)
PY
  done
done
