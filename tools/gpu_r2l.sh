#!/bin/bash
# stencil / cell-width sweep of the pair search on the plaza (per-row loops)
mkdir -p gpurun_out
for v in base xsub2 xsub8; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
  for st in 1 2 3; do
    python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 --stencil $st > gpurun_out/ab_r2l_${v}_st$st.json 2> gpurun_out/ab_r2l_${v}_st$st.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r2l_${v}_st$st.json')); print('$v', 'st$st', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()}, d['roofline']['fp32']['pair_tests_per_agent'])
except Exception as e: print('$v st$st ERR', e)
PY
  done
done
