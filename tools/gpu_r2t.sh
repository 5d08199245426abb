#!/bin/bash
# DRAM traffic of every kernel of the frame over steady-state frames (re-sort frames included), final code
mkdir -p gpurun_out
python bench.py --steps 8 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/bench_r2t_pre.json 2> gpurun_out/bench_r2t_pre.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_r2t_plaza1m.csv \
  python bench.py --steps 16 --warmup 16 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_launch_r2t.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 900 --csv --log-file gpurun_out/launches_r2t_plaza1m_warm.csv \
  python bench.py --steps 16 --warmup 16 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_launch_r2t_warm.log 2>&1; echo "ncu launches (warm caches) rc=$?"
for k in k_pairs k_force k_collide; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^$k" -s 12 -c 1 -f -o gpurun_out/prof_${k}_r2t \
  python bench.py --steps 4 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_${k}_r2t.log 2>&1; echo "ncu $k rc=$?"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_batch_rollout" -c 1 -f -o gpurun_out/prof_batch_r2t \
  python bench.py --workload pso4096 --steps 1 --warmup 0 --no-cpu --no-e2e > gpurun_out/ncu_batch_r2t.log 2>&1; echo "ncu batch rc=$?"
