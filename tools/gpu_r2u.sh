#!/bin/bash
mkdir -p gpurun_out
for v in base u4 u8 u16; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
  for w in pso512 pso4096; do
    python bench.py --no-cpu --no-e2e --workload $w --steps 3 --warmup 1 > gpurun_out/ab_r2u_${v}_$w.json 2> gpurun_out/ab_r2u_${v}_$w.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r2u_${v}_$w.json')); print('$v', '$w', '%.3e'%d['value'], '%.3f ms'%d['ms_per_step'], d.get('scenario_evals_per_sec') or d.get('secondary'))
except Exception as e: print('$v $w ERR', e)
PY
  done
done
