#!/bin/bash
# k_pairs: survivors queued per thread and recorded at the end (LC_NOTE_QUEUE = 4 default, 2, 8) vs recorded on the spot (0)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_walled.py tests/test_gpu_domain.py -x -q > gpurun_out/pytest_gpu_r2w.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2w.log
tail -4 gpurun_out/pytest_gpu_r2w.log
for v in base q0 q2 q8; do
  if [ $v = base ]; then unset LEARN_COMMS_LIB; else export LEARN_COMMS_LIB=$PWD/learn_comms_b200/liblearncomms_$v.so; fi
  for w in plaza1m corridor100k; do
    python bench.py --no-cpu --no-e2e --no-secondary --steps 64 --warmup 16 --workload $w > gpurun_out/ab_r2w_${v}_$w.json 2> gpurun_out/ab_r2w_${v}_$w.err
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/ab_r2w_${v}_$w.json')); print('$v', '$w', '%.3e'%d['value'], '%.4f'%d['ms_per_step'], d['state_hash']['hash'], {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms'].items()})
except Exception as e: print('$v $w ERR', e)
PY
  done
done
unset LEARN_COMMS_LIB
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^k_pairs" -s 12 -c 1 -f -o gpurun_out/prof_k_pairs_r2w \
  python bench.py --steps 4 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_k_pairs_r2w.log 2>&1; echo "ncu rc=$?"
