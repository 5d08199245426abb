"""The reference's own optimiser, UNMODIFIED (pso_opt.py, copied by __graft_entry__.build() from /root/reference into the
git-ignored oracle/_ref/), driving the drop-in executables through the process / file boundary it was written for
(pso_opt.py:236-267 simulate_set / simulate, :271-308 record / validate): 2 particles, 1 iteration, a wall-free and a
walled training scene, one validation scene.  ffmpeg (pso_opt.py:288-291 hstack of the two recordings) does not exist on
this box and there is no graphics stack to record with: a stub on PATH stands in for it, nothing else is touched."""
import os
import stat
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PSO = os.path.join(ROOT, "oracle", "_ref", "pso_opt.py")

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not os.path.exists(PSO), reason="oracle/_ref/pso_opt.py missing (built where /root/reference is present)")
def test_reference_pso_opt_runs_against_the_dropin(tmp_path):
    from learn_comms_b200 import build
    build.build()
    subprocess.check_call(["bash", os.path.join(ROOT, "tools", "build_cli.sh")], stdout=subprocess.DEVNULL)
    (tmp_path / "data").mkdir()
    stub = tmp_path / "stubs"
    stub.mkdir()
    ff = stub / "ffmpeg"
    ff.write_text("#!/bin/sh\nexit 0\n")
    ff.chmod(ff.stat().st_mode | stat.S_IEXEC)
    env = dict(os.environ, LEARN_COMMS_PROJECT_DIR=str(tmp_path), PATH=str(stub) + os.pathsep + os.environ["PATH"])
    cmd = [sys.executable, PSO, "run1", "--n_particles", "2", "--iters", "1", "--seed", "0", "--batch", "4", "13", ":", "4",
           "--program", os.path.join(ROOT, "learn_comms_b200")]
    r = subprocess.run(cmd, cwd=tmp_path, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    run = tmp_path / "data" / "run1"
    loss = (run / "loss.tsv").read_text().splitlines()
    assert loss[0].split("\t") == ["model#", "min p", "L(c)", "L(ttc)", "exit"]
    model, min_p, l_c, l_ttc, how = loss[1].split("\t")
    assert how == "iter" and np.isfinite(float(l_c)) and 0 < float(l_c) < 1e6
    assert float(l_ttc) == 2.0   # the baseline's loss normalised by itself, summed over the two training scenes
    # every evaluation left its archive: config + result per (particle, scene), the format pso_opt.py:253-265 writes
    for scene in (4, 13):
        it = run / "t0_train" / f"s{scene}_iters"
        names = sorted(p.name for p in it.iterdir() if p.is_dir())
        assert names == ["0", "1_p0", "1_p1", "2_p0", "2_p1"], names
        for nme in names:
            res = (it / nme / f"{nme}.result").read_text().split()
            assert len(res) == 2 and np.isfinite(float(res[1]))
    cv = (run / "t0_min" / "cross_val.tsv").read_text().splitlines()
    assert cv[0].startswith("model_id") and cv[1].split("\t")[0] == "4" and np.isfinite(float(cv[1].split("\t")[1]))
    assert "sums:" in r.stdout
