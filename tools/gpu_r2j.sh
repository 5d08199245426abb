#!/bin/bash
# Round 2 evidence run: full GPU suite, the default bench line (both arms), ncu launch list and --set full captures.
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2j.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2j.log
tail -8 gpurun_out/pytest_gpu_r2j.log
python bench.py > gpurun_out/bench_r2j_plaza1m.json 2> gpurun_out/bench_r2j_plaza1m.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r2j_reference.json 2> gpurun_out/bench_r2j_reference.err; echo "ref rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_r2j_plaza1m.json')); print(d['value'], d['ms_per_step'], d['e2e'], d['roofline'].get('kernel_ms'), d['cpu_baseline'])
d=json.load(open('gpurun_out/bench_r2j_reference.json')); print('ref', d['value'], d.get('cpu_baseline'))"
# launch list of the same command (cold-cache, serialised: shares only)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r2j_plaza1m.csv \
  python bench.py --steps 8 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_launch_r2j.log 2>&1; echo "ncu launches rc=$?"
for k in k_pairs k_force k_collide k_reorder; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^$k" -s 12 -c 1 -f -o gpurun_out/prof_${k}_r2j \
  python bench.py --steps 4 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_${k}_r2j.log 2>&1; echo "ncu $k rc=$?"
done
ls -la gpurun_out/*.ncu-rep
