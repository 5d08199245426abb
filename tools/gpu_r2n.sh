#!/bin/bash
mkdir -p gpurun_out
for v in base zc sc zcsc; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
    python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 > gpurun_out/ab_r2n_${v}.json 2> gpurun_out/ab_r2n_${v}.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r2n_${v}.json')); print('$v', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()}, d['timeline']['entry_to_entry'])
except Exception as e: print('$v ERR', e)
PY
done
unset LEARN_COMMS_LIB
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^k_force" -s 12 -c 1 -f -o gpurun_out/prof_k_force_r2n \
  python bench.py --steps 4 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_k_force_r2n.log 2>&1; echo "ncu rc=$?"
