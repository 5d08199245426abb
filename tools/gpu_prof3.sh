#!/bin/bash
# full ncu capture of the single-block kernels on the 125k plaza
mkdir -p gpurun_out
for k in k_frame_begin k_scan_hull k_bin; do
ncu --set full --clock-control none --import-source on -k $k -s 30 -c 1 -f -o gpurun_out/prof125_$k python bench.py --workload plaza125k --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu125_$k.log 2>&1
done
ls -la gpurun_out/prof125_*
