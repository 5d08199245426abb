#!/bin/bash
# full ncu capture of the two halves of the force pass (one launch each, frame 30 of the plaza run)
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/prof2_bench.json 2> gpurun_out/prof2_bench.err || exit 1
for k in k_force k_pairs; do
ncu --set full --clock-control none --import-source on -k $k -s 30 -c 1 -f -o gpurun_out/prof_$k python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
