"""CPU suite (-m "not gpu"): the oracle against known answers derived from the reference source, the host-side
logic, and the C ABI's export table.  No compute entry point of liblearncomms.so is called here (no GPU)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import _oracle
from learn_comms_b200 import workloads as wl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---------------------------------------------------------------- RNG (Seeder.cpp:5-20, LMP.cpp:261-263)
def test_lcg_first_draws():
    L = _oracle.lib()
    x = L.canon_lcg_seed(0)
    assert x == 1  # seed 0 -> state 1 (libstdc++ linear_congruential_engine::seed)
    seq = []
    for _ in range(3):
        x = L.canon_lcg_next(x)
        seq.append(x)
    assert seq == [16807, 282475249, 1622650073]  # minstd_rand0, SURVEY.md appendix A probe


def test_uniform_matches_libstdcxx(tmp_path):
    """canon_uniform against std::uniform_real_distribution<float>(-.01f,.01f)(std::default_random_engine) compiled here."""
    src = tmp_path / "probe.cpp"
    src.write_text(
        "#include <random>\n#include <cstdio>\n#include <cstring>\n"
        "int main(){std::default_random_engine g; g.seed(7); std::uniform_real_distribution<float> d(-.01f,.01f);"
        "for(int i=0;i<20000;++i){float v=d(g); unsigned u; memcpy(&u,&v,4); printf(\"%u\\n\",u);} }\n")
    exe = tmp_path / "probe"
    subprocess.check_call(["g++", "-O1", "-o", str(exe), str(src)])
    ref = np.array(subprocess.check_output([str(exe)]).split(), dtype=np.uint32)
    L = _oracle.lib()
    x = L.canon_lcg_seed(7)
    got = np.zeros(20000, np.float32)
    for i in range(20000):
        x = L.canon_lcg_next(x)
        got[i] = L.canon_uniform(x, -0.01, 0.01)
    assert np.array_equal(got.view(np.uint32), ref)


# ---------------------------------------------------------------- known-answer micro-cases
def two_agents(p0, p1, g0, g1, **kw):
    pos = np.array([p0, p1], np.float32)
    goal = np.array([g0, g1], np.float32)
    return _oracle.CanonSim(pos, goal, max_duration=45.0, rng_state=1, **kw)


def test_head_on_pair_ttc():
    """p = (-+1, 0), v = (+-1, 0): a = 4, b = -4, c = 3.91 -> t = (4 - sqrt(0.36)) / 4 = 0.85 (BoundVolume.cpp:102-122)."""
    o = two_agents((-1, 0), (1, 0), (10, 0), (-10, 0))
    st = o.state()
    st["vel"][:] = [[1, 0], [-1, 0]]
    st["vfh"][:] = [[1, 0], [-1, 0]]
    o.set_state(**st)
    o.record_neighbors(True)
    o.step(1)
    assert list(o.neighbor_counts()) == [1, 1]
    # force on agent 0: goal_F = 2*(goal_vel - vel) ~ 0, ttc_F = -x * (5 - .85)/(.85 + .001); vfh = 1 + dt*F -> clamped to unit
    s = o.state()
    t = np.float32((4 - np.sqrt(np.float32(16) - np.float32(4) * np.float32(3.91))) / 4)
    assert abs(float(t) - 0.85) < 1e-5
    mag = (5.0 - float(t)) / (float(t) + 0.001)
    expect_vfh0 = 1.0 + (1 / 60) * (-mag)
    assert abs(s["vfh"][0, 0] - expect_vfh0) < 1e-3
    assert s["vfh"][1, 0] == -s["vfh"][0, 0]  # Newton-3 symmetry is exact


def test_stationary_pair_no_interaction():
    """a = 0 -> t = NaN or inf -> FLT_MAX (BoundVolume.cpp:117, LMP.cpp:31-34): not a neighbour."""
    o = two_agents((-1, 0), (1, 0), (-1, 5), (1, 5))
    o.step(1)
    assert list(o.neighbor_counts()) == [0, 0]
    assert o.overlap_pairs == 0


def test_overlapping_pair_t0_mag20():
    o = two_agents((0, 0), (0.2, 0), (0, 10), (0.2, 10))
    st = o.state()
    st["vel"][:] = [[0.1, 0], [-0.1, 0]]
    o.set_state(**st)
    o.step(1)
    assert list(o.neighbor_counts()) == [1, 1]
    assert o.overlap_pairs == 1
    assert o.collision_hits == 2  # both agents query each other (physics.cpp:58-69)


# ---------------------------------------------------------------- static walls
def _wall_words(rect, a, b, r=0.15):
    rec = np.array([list(rect) + list(a) + list(b) + [r]], np.float32)
    w = _oracle.wall_geometry(rec)[0]
    return w, w.view(np.float32)


def test_wall_known_answers():
    """Hand-derived cases for a 1 m square wall tile at the origin (BoundVolume.cpp:131-188, LMP.cpp:64-82,307-319)."""
    tile = (0, 0, 1, 1)
    fmax = np.finfo(np.float32).max
    # agent 2 m to the left walking right at 1 m/s: the wall extruded by r starts at x = -0.65 -> t = 1.35 s
    w, f = _wall_words(tile, (-2, 0), (1, 0))
    assert abs(f[0] - 1.5) < 1e-6           # bare ray vs the tile face at x = -0.5
    assert abs(f[3] - 1.35) < 1e-6
    assert w[8] == 1 and f[9] < 0 and f[10] == 0  # pushes back along -x; mag = (5 - t)/(t + .001)
    assert abs(f[9] + (5 - 1.35) / 1.351) < 1e-4
    # walking away, or on a parallel track beside the tile: no time to collision, no force
    w, f = _wall_words(tile, (-2, 0), (-1, 0))
    assert f[3] == fmax and w[8] == 0
    w, f = _wall_words(tile, (-2, 0.7), (1, 0))
    assert f[0] == fmax and f[3] == fmax
    # ... but a track 0.6 m off-centre clips the r-inflated tile (0.5 + 0.15)
    w, f = _wall_words(tile, (-2, 0.6), (1, 0))
    assert f[0] == fmax and f[3] < 5
    # quirk Q10: the gate compares SQUARED distance with 5 (LMP.cpp:310), so a wall 2.3 m away is ignored
    w, f = _wall_words(tile, (-2.3, 0), (1, 0))
    assert f[3] < 5 and w[8] == 0
    # sight lines: through the tile blocked, past it free; the planner's c-space inflates by the agent's square
    w, _ = _wall_words(tile, (-2, 0), (2, 0))
    assert w[1] == 0 and w[7] == 0
    w, _ = _wall_words(tile, (-2, 0.6), (2, 0.6))
    assert w[1] == 1 and w[7] == 0          # clear of the tile, inside the inflated one (0.5 + 0.15)
    w, _ = _wall_words(tile, (-2, 0.7), (2, 0.7))
    assert w[1] == 1 and w[7] == 1
    # stationary agent: the slab test treats |v| < 1e-7 as parallel (BoundVolume.cpp:167-171)
    w, f = _wall_words(tile, (-2, 0), (0, 0))
    assert f[0] == fmax and f[3] == fmax


def test_wall_contact_normal():
    """BVH.cpp:121-130 + BoundVolume.cpp:196-230: an agent overlapping a face by 0.05 m is pushed out by exactly that."""
    w, f = _wall_words((0, 0, 1, 1), (-0.6, 0), (0, 0))
    assert w[4] == 1
    assert abs(f[5] + 0.05) < 1e-6 and f[6] == 0
    w, f = _wall_words((0, 0, 1, 1), (-0.66, 0), (0, 0))
    assert w[4] == 0
    # centre of the agent on the centre of the wall: normalize(0) -> NaN, as in the reference
    w, f = _wall_words((0, 0, 1, 1), (0, 0), (0, 0))
    assert np.isnan(f[5])


def test_device_wall_header_matches_reference_golden_when_compiled_for_the_host(tmp_path):
    """learn_comms_b200/csrc/walls.cuh (what k_wall_geometry runs) compiled by g++ with contraction off: every word
    equals the reference's, NaN patterns included.  The -m gpu suite repeats this on the device through the C ABI."""
    so = tmp_path / "walls_host.so"
    subprocess.check_call(["g++", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math", "-w", "-x", "c++",
                           os.path.join(ROOT, "tests", "walls_host_shim.cpp"), "-o", str(so)])
    L = C.CDLL(str(so))
    g = np.load(os.path.join(ROOT, "tests", "golden", "wall_geometry.npz"))
    rec = np.ascontiguousarray(g["records"], np.float32)
    out = np.zeros((len(rec), 12), np.uint32)
    L.walls_host_batch(C.c_int32(len(rec)), rec.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    assert np.array_equal(out, g["words"])
    assert np.array_equal(out, _oracle.wall_geometry(rec))


def test_walled_rollouts_are_refused_by_the_oracle():
    """Rollouts with walls need the PRM + A* planner (GMP.cpp), which is not restated: canon_create says no."""
    L = _oracle.lib()
    pos = np.zeros((1, 2), np.float32)
    wall = _oracle.Rect(0, 0, 1, 1)
    d = _oracle.Desc(1, pos.ctypes.data, pos.ctypes.data, 0.15, 45.0, 1, None, None, 1, 0, 1, C.addressof(wall))
    assert L.canon_create(C.byref(d)) is None


def test_zero_config_message_is_half_and_no_comm_force():
    """All-zero comms.config == pure-TTC baseline: c == 0.5 after frame 1 (CommComp.h:20-27), comm force == 0."""
    pos, goal, md, rng = _oracle.scene(0, 0)
    a = _oracle.CanonSim(pos, goal, None, None, max_duration=md, rng_state=rng, comm=True)
    b = _oracle.CanonSim(pos, goal, None, None, max_duration=md, rng_state=rng, comm=False)
    a.step(200)
    b.step(200)
    sa, sb = a.state(), b.state()
    assert (sa["c"] == 0.5).all()
    for k in ("pos", "vel", "vfh"):
        assert np.array_equal(sa[k], sb[k])


def test_single_path_done_within_1m_and_min_time():
    """Wall-free: plan = [goal]; done when within 1 m (LMP.cpp:87); min_time = |goal - start| - 1 (ai.cpp:121-126)."""
    o = two_agents((0, 0), (100, 100), (5, 0), (100, 105))
    k = o.step(10 ** 5)
    sc = o.scores()
    assert o.finished and sc["n_done"] == 2
    # unit speed after a short ramp: both walk 4 m; overhead = arrival time - 4 s is a fraction of a second
    assert 0.0 < sc["run_avg"] < 1.0
    assert 4.0 * 60 < k < 5.0 * 60 + 5


def test_timeout_penalty():
    """demo.cpp:691-700: not-done agents contribute 2*|goal - pos| + t - min_time."""
    o2 = _oracle.CanonSim(np.array([[0, 0], [50, 50]], np.float32), np.array([[1000, 0], [50, 1050]], np.float32),
                          max_duration=1.0, rng_state=1)
    k = o2.step(1000)
    assert k == 61  # first frame with dt*frame > 1.0
    sc = o2.scores()
    # walked ~0.85 m of 1000: 2*999.15 + 1.0167 - (1000 - 1) ~ 1000.3 .. 1001
    assert sc["n_done"] == 2 and 1000.0 < sc["run_avg"] < 1001.5


def test_farthest_partner_quirk_q1():
    """select_nearest returns the FARTHEST agent (comm.cpp:204,214-217)."""
    pos = np.array([[0, 0], [1, 0], [5, 0], [-7, 0]], np.float32)
    o = _oracle.CanonSim(pos, pos + 1, max_duration=45.0, rng_state=1)
    o.step(1)
    assert list(o.partners()) == [3, 3, 3, 2]


# ---------------------------------------------------------------- grid search == brute force inside the oracle
@pytest.mark.parametrize("seed", [0, 1])
def test_oracle_grid_equals_bruteforce(seed):
    pos, goal, md, rng = wl.plaza(3000, 110.0, seed=seed, goal_range=25.0)
    m_cv, m_f = wl.weights(seed + 40)
    a = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, grid=False)
    b = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, grid=True)
    a.record_neighbors(True)
    b.record_neighbors(True)
    for f in range(25):
        a.step(1)
        b.step(1)
        sa, sb = a.state(), b.state()
        for k in sa:
            assert np.array_equal(sa[k], sb[k]), (f, k)
        assert np.array_equal(a.partners(), b.partners())
        pa, ia = a.neighbor_lists()
        pb, ib = b.neighbor_lists()
        assert np.array_equal(pa, pb) and np.array_equal(ia, ib)
        assert a.overlap_pairs == b.overlap_pairs and a.collision_hits == b.collision_hits


def test_oracle_grid_equals_bruteforce_dense_from_rest():
    """Packed crowd starting at rest: v_max = 0 makes the TTC cutoff (0.31 m) smaller than the contact reach of the
    collision pass (0.3 + 2 dt); the grid must still find every pair that overlaps after the move."""
    rs = np.random.RandomState(4)
    n = 600
    pos = rs.uniform(-3, 3, (n, 2)).astype(np.float32)
    goal = rs.uniform(-3, 3, (n, 2)).astype(np.float32)
    m_cv, m_f = wl.weights(8)
    a = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=30.0, rng_state=1, grid=False)
    b = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=30.0, rng_state=1, grid=True)
    hits = 0
    for f in range(12):
        a.step(1)
        b.step(1)
        sa, sb = a.state(), b.state()
        for k in sa:
            assert np.array_equal(sa[k], sb[k], equal_nan=True), (f, k)
        assert a.overlap_pairs == b.overlap_pairs and a.collision_hits == b.collision_hits
        hits += a.collision_hits
    assert hits > 0


def test_neighbor_predicate_symmetry():
    """ttc(i,j) == ttc(j,i) bit for bit: neighbour lists are symmetric."""
    pos, goal, md, rng = _oracle.scene(12, 0)
    o = _oracle.CanonSim(pos, goal, *wl.weights(3), max_duration=md, rng_state=rng)
    o.record_neighbors(True)
    o.step(120)
    ptr, idx = o.neighbor_lists()
    pairs = {(i, int(j)) for i in range(o.n) for j in idx[ptr[i]:ptr[i + 1]]}
    assert pairs and all((j, i) in pairs for (i, j) in pairs)


# ---------------------------------------------------------------- files (comm.cpp:74-142, pso_opt.py:17-41)
def test_config_roundtrip_pso_format(tmp_path):
    cfg = np.random.RandomState(0).uniform(-1, 1, (5, 8))
    cfg[3:, 3:] = 0
    p = tmp_path / "comms.config"
    with open(p, "w") as f:  # pso_opt.py:17-23 write_config
        for row in cfg:
            f.write(" ".join(str(v) for v in row) + "\n")
    m_cv = np.zeros((3, 8), np.float32)
    m_f = np.zeros((2, 3), np.float32)
    assert _oracle.lib().canon_read_config(os.fsencode(p), _oracle._p(m_cv), _oracle._p(m_f)) == 0
    assert np.array_equal(m_cv, cfg[:3].astype(np.float32))
    assert np.array_equal(m_f, cfg[3:, :3].astype(np.float32))


def test_legacy_config_is_rejected(tmp_path):
    """HEAD's parser aborts on the legacy '3 5'-headed files: stof("") on the missing fifth line."""
    p = tmp_path / "legacy.config"
    p.write_text("3 5\n0.1 0.2 0.3 0.4 0.5\n0 0 0 0 0\n0.3 0 0 0 0\n")
    m_cv = np.zeros((3, 8), np.float32)
    m_f = np.zeros((2, 3), np.float32)
    assert _oracle.lib().canon_read_config(os.fsencode(p), _oracle._p(m_cv), _oracle._p(m_f)) != 0


def test_result_format(tmp_path):
    p = tmp_path / "comms.result"
    assert _oracle.lib().canon_write_result(os.fsencode(p), 0.404623, 120.279) == 0
    assert p.read_text() == "0.404623\n120.279\n"  # comms_82.result:1-2 formatting (6 significant digits)


# ---------------------------------------------------------------- the C ABI library: loads and exports every declared symbol
def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "learn_comms.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(lc_[a-z_0-9]+)\s*\(", hdr)))


def test_cabi_exports_every_declared_symbol():
    from learn_comms_b200 import build
    so = build.build()
    L = C.CDLL(so)
    names = _declared_symbols()
    assert len(names) >= 29
    for nme in names:
        assert hasattr(L, nme), f"{nme} declared in include/learn_comms.h but not exported"
    L.lc_version.restype = C.c_char_p
    assert b"sm_100a" in L.lc_version()


def test_host_files_and_scenes_match_oracle(tmp_path):
    """The product's host-side (non-GPU) entry points against the oracle: scene catalogue, config parser, result writer."""
    import learn_comms_b200 as lcb
    for sc in (0, 1, 4, 5, 12):
        for seed in (0, 3):
            a, b = lcb.scene(sc, seed), _oracle.scene(sc, seed)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2:] == b[2:]
    p = tmp_path / "comms.config"
    cfg = np.random.RandomState(1).uniform(-1, 1, (5, 8))
    with open(p, "w") as f:
        for row in cfg:
            f.write(" ".join(str(v) for v in row) + "\n")
    m_cv, m_f = lcb.read_config(str(p))
    assert np.array_equal(m_cv, cfg[:3].astype(np.float32)) and np.array_equal(m_f, cfg[3:, :3].astype(np.float32))
    with pytest.raises(lcb.LcError):
        lcb.read_config(str(tmp_path / "missing.config"))
    r = tmp_path / "comms.result"
    lcb.write_result(str(r), 0.404623, 120.279)
    assert r.read_text() == "0.404623\n120.279\n"


def test_no_device_fails_loudly():
    """No CPU fallback: without a CUDA device lc_create must return LC_ERR_CUDA (skipped where a GPU exists)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import learn_comms_b200 as lcb
    pos, goal, md, rng = lcb.scene(0, 0)
    with pytest.raises(lcb.LcError) as e:
        lcb.Sim(pos, goal)
    assert e.value.code == -2


def test_no_packed_fma_in_sass():
    """ptxas contracts packed mul + add into FFMA2 even under -fmad=false (single rounding -> neighbour lists would
    differ from the reference's SSE2 mul/add).  The kernels avoid the pattern; make sure no FFMA2 ever reappears."""
    import shutil
    from learn_comms_b200 import build
    so = build.build()
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.check_output([cuobjdump, "-sass", so], text=True)
    assert "FMUL2" in sass and "FADD2" in sass  # the packed pair loop is there
    assert "FFMA2" not in sass
