#!/bin/bash
mkdir -p gpurun_out
for v in base cb12 cb16; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
  python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 > gpurun_out/ab_r3n_${v}.json 2> gpurun_out/ab_r3n_${v}.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r3n_${v}.json')); print('$v', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()})
except Exception as e: print('$v ERR', e)
PY
done
