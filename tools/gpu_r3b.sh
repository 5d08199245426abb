#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^k_force" -s 12 -c 1 -f -o gpurun_out/prof_k_force_r3b \
  python bench.py --steps 4 --warmup 3 --no-cpu --no-secondary --no-e2e > gpurun_out/ncu_k_force_r3b.log 2>&1; echo "ncu rc=$?"
