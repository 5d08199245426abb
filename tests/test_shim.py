"""Source-level drop-in, COMPILED AND RUN: oracle/_ref/oref_gpu_shim = the reference engine's own
translation units (demo, comm, ai, physics, Pool: unmodified, from /root/reference) with the per-frame hot path of
main.cpp:131-144 replaced by liblearncomms.so (tests/shim/gpu_path.cpp).  Scene construction, the first plans, done
detection, Welford statistics, the time-out penalty and the result file are the reference's own code running on the
state the GPU produced.  Held against (a) the process-level drop-in (bin/comm_norender) — same comms.result, same
per-finisher log — and (b) the reference's own result files (tests/golden, from O-ref) on rollouts in which the
reference's Gauss-Seidel collision order never mattered."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "oracle", "_ref", "oref_gpu_shim")
BIN = os.path.join(ROOT, "learn_comms_b200", "bin")
GOLD = os.path.join(ROOT, "tests", "golden")

pytestmark = pytest.mark.gpu


def _config(d, wseed):
    from learn_comms_b200 import pso, workloads as wl
    cfg = np.zeros(pso.M_SHAPE)
    if wseed is not None:
        m_cv, m_f = wl.weights(wseed)
        cfg[:3] = m_cv
        cfg[3:, :3] = m_f
    with open(os.path.join(d, "comms.config"), "w") as f:
        pso.write_config(f, cfg)


@pytest.mark.skipif(not os.path.exists(SHIM), reason="oracle/_ref/oref_gpu_shim not built (needs /root/reference at build time)")
@pytest.mark.parametrize("scene,wseed,golden", [(4, 2, None), (18, None, "walled_rollout_18_zero"), (10, 30, "walled_rollout_10_w30"),
                                                (13, 33, "walled_rollout_13_w33")])
def test_reference_engine_with_gpu_hot_path(tmp_path, scene, wseed, golden):
    subprocess.check_call(["bash", os.path.join(ROOT, "tools", "build_cli.sh")], stdout=subprocess.DEVNULL)
    a = tmp_path / "shim"
    b = tmp_path / "data" / "cli"
    a.mkdir()
    b.mkdir(parents=True)
    if golden is None:
        _config(str(a), wseed)
    else:  # the very config text the fixture's reference run read (tests/_oref.py write_config)
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import _oref
        g = np.load(os.path.join(GOLD, golden + ".npz"))
        _oref.write_config(str(a / "comms.config"), g["m_cv"] if "m_cv" in g else None, g["m_f"] if "m_f" in g else None)
    (b / "comms.config").write_text((a / "comms.config").read_text())
    r1 = subprocess.run([SHIM, str(scene), "0", str(a)], capture_output=True, text=True, timeout=900)
    assert r1.returncode == 0, r1.stderr[-2000:]
    env = dict(os.environ, LEARN_COMMS_PROJECT_DIR=str(tmp_path))
    r2 = subprocess.run([os.path.join(BIN, "comm_norender"), str(scene), "--seed", "0", "--data", "cli"], capture_output=True, text=True, env=env, timeout=900)
    assert r2.returncode == 0, r2.stderr[-2000:]
    res_shim, res_cli = (a / "comms.result").read_text(), (b / "comms.result").read_text()
    assert res_shim == res_cli, (res_shim, res_cli)
    # demo.cpp:662: the reference's own update_runs printed one line per finisher; the drop-in prints the same lines
    t1 = [l for l in r1.stdout.splitlines() if l.startswith("t:")]
    t2 = [l for l in r2.stdout.splitlines() if l.startswith("t:")]
    assert t1 == t2 and len(t1) > 0
    assert "FRAMES: " in r1.stderr and r1.stderr.split("FRAMES: ")[1].split()[0] == r2.stderr.split("FRAMES: ")[1].split()[0]
    if golden is not None:
        assert tuple(res_shim.split()) == tuple(str(x) for x in g["result_file"]), "differs from the reference's own comms.result"
        assert int(r1.stderr.split("FRAMES: ")[1].split()[0]) == int(g["frames_executed"])
