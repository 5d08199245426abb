#!/bin/bash
# A/B of differently built k_force variants on the same box (LEARN_COMMS_LIB override)
mkdir -p gpurun_out
for v in base pf8 np8 pf6; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
  for w in plaza1m plaza125k; do
    python bench.py --no-cpu --no-e2e --workload $w > gpurun_out/ab_${v}_$w.json 2> gpurun_out/ab_${v}_$w.err
  done
done
python - <<'PY'
import json
for v in ('base','pf8','np8','pf6'):
    for w in ('plaza1m','plaza125k'):
        try:
            d=json.load(open(f'gpurun_out/ab_{v}_{w}.json')); print(v, w, '%.3e'%d['value'], '%.4f'%d['ms_per_step'], round(d['roofline']['kernel_ms']['k_force']*1e3,1))
        except Exception as e: print(v,w,'ERR',e)
PY
