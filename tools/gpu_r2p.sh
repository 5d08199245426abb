#!/bin/bash
# 2-GPU box: IPC-path domain test, then plaza-1M at N = 1 and N = 2, frozen ghost sets (re-sort every 8) vs routing every frame
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_domain_ipc.py tests/test_gpu_domain.py -x -q > gpurun_out/pytest_gpu_r2p_domain_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_r2p_domain_2gpu.log
tail -4 gpurun_out/pytest_gpu_r2p_domain_2gpu.log
EXTRA="--no-cpu --no-secondary --no-e2e"
python bench.py --steps 64 --warmup 16 $EXTRA > gpurun_out/bench_r2p_n1.json 2> gpurun_out/bench_r2p_n1.err; echo "n1 rc=$?"
for k in 8 1 16; do
LC_REBIN_EVERY=$k timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$k bench.py --gpus 2 --steps 64 --warmup 16 $EXTRA > gpurun_out/bench_r2p_n2_k$k.json 2> gpurun_out/bench_r2p_n2_k$k.err; echo "n2 k=$k rc=$?"; tail -2 gpurun_out/bench_r2p_n2_k$k.err
done
python - <<'PY'
import json
for f in ("n1", "n2_k8", "n2_k1", "n2_k16"):
    try:
        d = json.load(open(f"gpurun_out/bench_r2p_{f}.json"))
        print(f, '%.3e' % d["value"], '%.1f us' % (d["ms_per_step"] * 1e3), d["state_hash"]["hash"], {k: round(v * 1e3, 1) for k, v in d["roofline"]["kernel_ms"].items()}, d.get("timeline"))
    except Exception as e:
        print(f, "failed", e)
PY
