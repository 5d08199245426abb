#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2h.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_r2h.log
tail -12 gpurun_out/pytest_gpu_r2h.log
for k in 1 8; do
LC_REBIN_EVERY=$k python bench.py --steps 64 --warmup 16 --no-cpu --no-secondary --no-e2e > gpurun_out/bench_r2h_k$k.json 2> gpurun_out/bench_r2h_k$k.err
python -c "
import json; d=json.load(open('gpurun_out/bench_r2h_k$k.json')); print('K=$k', d['value'], d['ms_per_step'], d['state_hash']['hash'], d['timeline']['entry_to_entry'], d['roofline']['kernel_ms'])"
done
python bench.py --steps 20 --warmup 5 --no-cpu --no-secondary --no-e2e > gpurun_out/bench_r2h_s20.json 2> gpurun_out/bench_r2h_s20.err
python -c "
import json; d=json.load(open('gpurun_out/bench_r2h_s20.json')); print('steps20', d['value'], d['ms_per_step'], d['state_hash']['hash'])"
