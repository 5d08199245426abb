#!/bin/bash
# Round 2, first GPU visit: whole -m gpu suite, smoke, default bench line at the driver's step counts, reference arm.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2a.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2a.log
tail -15 gpurun_out/pytest_gpu_r2a.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_r2a.log 2>&1; tail -2 gpurun_out/smoke_r2a.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r2a_plaza1m.json 2> gpurun_out/bench_r2a_plaza1m.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_r2a_plaza1m.json; tail -3 gpurun_out/bench_r2a_plaza1m.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_r2a_reference.json 2> gpurun_out/bench_r2a_reference.err; tail -c 1200 gpurun_out/bench_r2a_reference.json
nproc
