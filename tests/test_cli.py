"""The comm_norender / nocomm_norender drop-in executables (process/file boundary of pso_opt.py:245-267)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "learn_comms_b200", "bin")


def _build():
    from learn_comms_b200 import build
    build.build()
    subprocess.check_call(["bash", os.path.join(ROOT, "tools", "build_cli.sh")], stdout=subprocess.DEVNULL)


def _run(exe, args, project):
    env = dict(os.environ, LEARN_COMMS_PROJECT_DIR=str(project))
    return subprocess.run([os.path.join(BIN, exe)] + [str(a) for a in args], capture_output=True, text=True, env=env, timeout=600)


def test_cli_argument_grammar(tmp_path):
    """io.cpp:63-125: misuse -> message + usage on stderr, exit status 1; --help -> usage on stdout, exit 0."""
    _build()
    r = _run("comm_norender", [], tmp_path)
    assert r.returncode == 1 and "missing required positional" in r.stderr and "Usage:" in r.stderr
    r = _run("comm_norender", ["--help"], tmp_path)
    assert r.returncode == 0 and r.stdout.startswith("Usage:") and "[--seed u]" in r.stdout and r.stdout.rstrip().endswith("scene")
    assert _run("comm_norender", [0, "--bogus"], tmp_path).returncode == 1
    assert "no double options" in _run("comm_norender", [0, "--seed", 1, "--seed", 2], tmp_path).stderr
    assert "no arg to read" in _run("comm_norender", [0, "--seed"], tmp_path).stderr
    assert "bad arg" in _run("comm_norender", [0, "--seed", "-3"], tmp_path).stderr
    assert "no further positionals" in _run("comm_norender", [0, 1], tmp_path).stderr
    # missing config: io.cpp:130 message, failure status
    r = _run("comm_norender", [0, "--seed", 0, "--data", "none"], tmp_path)
    assert r.returncode == 1 and "Failed to open file" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("exe,scene,seed", [("comm_norender", 4, 0), ("nocomm_norender", 0, 2), ("comm_norender", 12, 7),
                                            ("comm_norender", 18, 0), ("comm_norender", 8, 3), ("nocomm_norender", 13, 1)])
def test_cli_result_file_matches_oracle(tmp_path, exe, scene, seed):
    """Whole-process drop-in: config file in, comms.result out, byte-identical to what the oracle's writer produces."""
    import _oracle
    from learn_comms_b200 import pso, workloads as wl
    _build()
    d = tmp_path / "data" / "run" / "iters"
    d.mkdir(parents=True)
    cfg = np.zeros(pso.M_SHAPE)
    m_cv, m_f = wl.weights(scene + 10)
    cfg[:3] = m_cv
    cfg[3:, :3] = m_f
    with open(d / "comms.config", "w") as f:
        pso.write_config(f, cfg)
    r = _run(exe, [scene, "--seed", seed, "--data", "run/iters/"], tmp_path)
    assert r.returncode == 0, r.stderr
    pos, goal, md, rng, walls = _oracle.scene_full(scene, seed)
    # str(float64) -> stof: what both parsers see
    rc_cv, rc_f = np.zeros((3, 8), np.float32), np.zeros((2, 3), np.float32)
    assert _oracle.lib().canon_read_config(os.fsencode(str(d / "comms.config")), _oracle._p(rc_cv), _oracle._p(rc_f)) == 0
    import learn_comms_b200 as lcb
    order, wall_order = lcb.table_order(len(pos), len(walls))   # the drop-in runs in the reference's table iteration order (quirk Q4)

    def oracle():
        kw = dict(walls=walls, wall_order=wall_order, planner=lcb.Planner(walls, 0.15), bvh=lcb.static_bvh(walls, wall_order)) if len(walls) else {}
        return _oracle.CanonSim(pos, goal, rc_cv, rc_f, max_duration=md, rng_state=rng, comm=exe == "comm_norender", order=order, **kw)

    o = oracle()
    o.step(10 ** 6)
    sc = o.scores()
    ref = tmp_path / "oracle.result"
    _oracle.lib().canon_write_result(os.fsencode(str(ref)), sc["avg_velocity"], sc["group_time"])
    assert (d / "comms.result").read_text() == ref.read_text()
    assert pso.read_result(str(d / "comms.result")) == pso.read_result(str(ref))
    # agent_paths.csv (demo.cpp:678-684): n rows per executed frame — id, group, pos, vel, radius, time stamp — the way
    # fix_paths.py:15-25 reads them; the first frame's rows are the initial state, the last frame's the oracle's state
    # before its final frame
    rows = np.loadtxt(d / "agent_paths.csv", delimiter=",", dtype=np.float64)
    n = pos.shape[0]
    assert rows.shape == (o.frame * n, 8)
    assert np.array_equal(rows[:n, 0], order + 1) and not rows[:, 1].any()   # entity ids in POOL.for_<Agent> order
    assert np.allclose(rows[:n, 2:4], pos[order], rtol=1e-5, atol=1e-6) and not rows[:n, 4:6].any()
    assert np.allclose(rows[:, 6], 0.15) and rows[0, 7] == 0.0
    assert np.allclose(rows[-1, 7], (o.frame - 1) / 60.0, rtol=1e-5)
    # demo.cpp:662: one "t:%f u:%f s:%f" line per counted finisher on stdout, the last one carrying the final statistics
    lines = [l for l in r.stdout.splitlines() if l.startswith("t:")]
    assert len(lines) == sc["n_done"] and lines[-1].endswith("u:%f s:%f" % (sc["run_avg"], sc["run_std"]))
    o2 = oracle()
    o2.step(o.frame - 1)
    st = o2.state()
    assert np.allclose(rows[-n:, 2:4], st["pos"][order], rtol=1e-5, atol=1e-6) and np.allclose(rows[-n:, 4:6], st["vel"][order], rtol=1e-5, atol=1e-6)
    # same scores when the CSV is switched off (one lc_run_to_completion call instead of frame-by-frame stepping)
    env = dict(os.environ, LEARN_COMMS_PROJECT_DIR=str(tmp_path), LEARN_COMMS_NO_PATHS="1")
    first = (d / "comms.result").read_text()
    r2 = subprocess.run([os.path.join(BIN, exe), str(scene), "--seed", str(seed), "--data", "run/iters/"], capture_output=True, text=True, env=env, timeout=600)
    assert r2.returncode == 0 and (d / "comms.result").read_text() == first
