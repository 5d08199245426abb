#!/bin/bash
# One GPU-box visit: smoke, parity tests, both bench workloads + reference arm, ncu launch lists + one full capture per top kernel.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_plaza.json 2> gpurun_out/bench_plaza.err; tail -c 600 gpurun_out/bench_plaza.json
python bench.py --workload pso4096 --steps 5 --warmup 3 > gpurun_out/bench_pso.json 2> gpurun_out/bench_pso.err; cat gpurun_out/bench_pso.json; tail -3 gpurun_out/bench_pso.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
for w in corridor100k circle1k; do python bench.py --workload $w --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_plaza.csv python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_plaza.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 50 --csv --log-file gpurun_out/launches_pso.csv python bench.py --workload pso4096 --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_pso.log 2>&1
ncu --set full --clock-control none --import-source on -k k_batch_rollout -c 1 -f -o gpurun_out/prof_batch python bench.py --workload pso4096 --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_batch_full.log 2>&1
for k in k_force k_pairs; do
ncu --set full --clock-control none --import-source on -k $k -s 30 -c 1 -f -o gpurun_out/prof_$k python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_${k}_full.log 2>&1
done
ls -la gpurun_out | tail -30
