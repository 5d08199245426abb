#!/bin/bash
# Device parity of the static-wall geometry without pytest / torch start-up cost (fits a very short GPU slot).
mkdir -p gpurun_out
python - > gpurun_out/walls_gpu.log 2>&1 <<'PY'
import importlib.util, sys, time
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
spec = importlib.util.spec_from_file_location("t", "tests/test_gpu_walls.py")
t = importlib.util.module_from_spec(spec); spec.loader.exec_module(t)
for name in ("test_wall_geometry_matches_reference_golden", "test_wall_geometry_edges_and_errors",
             "test_wall_geometry_matches_oracle_on_fresh_records", "test_lookahead_matches_reference_golden",
             "test_lookahead_errors_and_no_walls"):
    t0 = time.time(); getattr(t, name)(); print(name, "PASSED", round(time.time() - t0, 2), "s", flush=True)
import __graft_entry__ as ge
t0 = time.time(); ge.smoke(); print("smoke() PASSED", round(time.time() - t0, 2), "s", flush=True)
PY
cat gpurun_out/walls_gpu.log
