#!/bin/bash
# final 1-GPU evidence: smoke, full GPU suite, default bench line (both arms), launch list
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_r3d.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_r3d.log
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r3d.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r3d.log
tail -4 gpurun_out/pytest_gpu_r3d.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r3d_reference.json 2> gpurun_out/bench_r3d_reference.err; echo "ref rc=$?"
python bench.py > gpurun_out/bench_r3d_plaza1m.json 2> gpurun_out/bench_r3d_plaza1m.err; echo "bench rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r3d_plaza1m_s20.json 2> gpurun_out/bench_r3d_plaza1m_s20.err; echo "bench s20 rc=$?"
python -c "
import json
for f in ('gpurun_out/bench_r3d_plaza1m.json','gpurun_out/bench_r3d_plaza1m_s20.json'):
    d=json.load(open(f)); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['dependent']['value'], d['e2e_read']['value'], d['roofline']['frac'], d['roofline']['frame'], d['roofline']['fp32']['achieved'], d['secondary']['scenario_evals_per_sec'], d['cpu_baseline']['value'], d['gpu_launches'], d['clocks'])
d=json.load(open('gpurun_out/bench_r3d_reference.json')); print('ref', d['value'], d['cpu_baseline']['cores'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r3d_plaza1m.csv \
  python bench.py --steps 8 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_launch_r3d.log 2>&1; echo "ncu launches rc=$?"
for k in k_pairs k_force; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^$k" -s 12 -c 1 -f -o gpurun_out/prof_${k}_r3d \
  python bench.py --steps 4 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_${k}_r3d.log 2>&1; echo "ncu $k rc=$?"
done
