"""Walled scenes on the GPU: lc_create with n_walls > 0 — per-agent
waypoint lists, the sight-line / lost_frames replan gate (ai.cpp:152-168) with replans served by the host planner between
frames, multi-waypoint lookahead (LMP.cpp:84-113), wall TTC forces (LMP.cpp:306-319) inside k_force, static-BVH push-out
(physics.cpp:60-69, BVH.cpp:106-138) inside k_collide.

Two bars, both bit for bit (0 ulp):
  1. against O-canon (Jacobi collision pass) in lock step, ascending and reference table order;
  2. against the REFERENCE'S OWN rollouts (tests/golden/*.npz from O-ref) wherever the reference's Gauss-Seidel pass and
     dynamic-BVH false negatives never came into play — whole 1 874- and 2 774-frame rollouts with dozens of replans."""
import os

import numpy as np
import pytest

import _oracle
import learn_comms_b200 as lcb
from learn_comms_b200 import workloads as wl

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KEYS = ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")


def _same(a, b, tag):
    for k in KEYS:
        same = np.array_equal(a[k], b[k], equal_nan=True) if a[k].dtype.kind == "f" else np.array_equal(a[k], b[k])
        if not same:
            bad = np.flatnonzero((a[k] != b[k]).reshape(len(a[k]), -1).any(1))
            raise AssertionError(f"{tag}: {k} differs for {bad.size} agents, first {bad[:6]}")


def _pair(scene, seed, wseed, reference_order, comm=True):
    pos, goal, md, rng, walls = lcb.scene_full(scene, seed)
    m_cv, m_f = (None, None) if wseed is None else wl.weights(wseed)
    order, wall_order = lcb.table_order(len(pos), len(walls)) if reference_order else (None, None)
    g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, comm=comm, walls=walls, order=order, wall_order=wall_order)
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, comm=comm, walls=walls, order=order, wall_order=wall_order,
                         planner=lcb.Planner(walls, 0.15), bvh=lcb.static_bvh(walls, wall_order))
    return g, o


@pytest.mark.parametrize("scene,wseed,ref_order,stride", [(3, 23, True, 9), (6, None, True, 7), (9, 29, True, 31), (13, 33, False, 5),
                                                          (18, 38, True, 13), (2, 22, False, 17), (11, 31, True, 29), (7, 27, True, 3),
                                                          (905, 5, True, 37)])
def test_walled_rollout_lockstep_vs_canon(scene, wseed, ref_order, stride):
    g, o = _pair(scene, 0, wseed, ref_order)
    f = 0
    while not o.finished:
        g.step(stride)
        o.step(stride)
        f += stride
        _same(g.state(), o.state(), f"scene {scene} frame {f}")
        wg, wo = g.walled(), o.walled()
        assert np.array_equal(wg["n_plan"], wo["n_plan"]) and np.array_equal(wg["lost_frames"], wo["lost_frames"]), f"frame {f}: plans / lost_frames"
        assert (wg["replans"], wg["wall_hits"]) == (wo["replans"], wo["wall_hits"]), f"frame {f}: {wg['replans'], wg['wall_hits']} vs {wo['replans'], wo['wall_hits']}"
        assert np.array_equal(g.partners(), o.partners()) and np.array_equal(g.neighbor_counts(), o.neighbor_counts())
        assert g.rng == o.rng
    assert g.finished and g.frame == o.frame
    gs, os_ = g.scores(), o.scores()
    for k in gs:
        assert gs[k] == os_[k] or (np.isnan(gs[k]) and np.isnan(os_[k])), (k, gs, os_)
    for i in (0, g.n // 2, g.n - 1):
        assert np.array_equal(g.plan(i), o.plan(i))


def test_walled_run_to_completion_and_set_state():
    """lc_run_to_completion over a walled scene (every pause served inside the call) and a get_state / set_state round
    trip in the middle (num_done is the waypoint counter here, the replan gate is re-evaluated on the forced state)."""
    g, o = _pair(10, 0, 30, True)
    g.step(700)
    o.step(700)
    st = g.state()
    assert st["num_done"].max() >= 1
    g.set_state(**st)
    g.step(100)
    o.step(100)
    _same(g.state(), o.state(), "after the round trip")
    frames = g.run_to_completion()
    o.step(10 ** 6)
    assert frames == o.frame == 2774 and g.scores() == o.scores()
    log = g.finish_log()   # demo.cpp:662: one (t, u, s) line per finisher
    assert len(log) == g.scores()["n_done"] and log[-1][1] == np.float32(g.scores()["run_avg"]) and log[-1][2] == np.float32(g.scores()["run_std"])


def test_walled_unreachable_goal_has_no_plan():
    """A goal inside a ring of walls: GMP::plan_one returns an empty path, has_plan() is false (Agent.h:20-22): the agent
    requests a replan every frame, draws no random number (LMP.cpp:267-270 — the others' draws shift), gets goal_vel = 0,
    and is scored through the time-out penalty."""
    rs = np.random.RandomState(5)
    ring = [[x, y, 2.0, 2.0] for x in (-2.0, 0.0, 2.0) for y in (-2.0, 0.0, 2.0) if (x, y) != (0.0, 0.0)]
    walls = np.array(ring, np.float32)
    pos = np.array([[-8, 0.2], [8, -0.3], [-8, 3], [8, 3.5], [0.2, 8], [0, -8]], np.float32)
    goal = np.array([[8, 0], [0.0, 0.0], [8, 3], [-8, 3.5], [0, -8], [0.3, 8]], np.float32)   # agent 1's goal is walled in
    m_cv, m_f = wl.weights(3)
    order, wall_order = lcb.table_order(len(pos), len(walls))
    g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=20.0, rng_state=77, walls=walls, order=order, wall_order=wall_order)
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=20.0, rng_state=77, walls=walls, order=order, wall_order=wall_order,
                         planner=lcb.Planner(walls, 0.15), bvh=lcb.static_bvh(walls, wall_order))
    assert o.walled()["n_plan"][1] == 0
    for f in range(0, 1300, 50):
        g.step(50)
        o.step(50)
        _same(g.state(), o.state(), f"frame {f + 50}")
        assert g.rng == o.rng
        if o.finished:
            break
    assert g.finished and o.finished and g.frame == o.frame and g.scores() == o.scores()
    assert g.walled()["replans"] == o.walled()["replans"] > 1000


@pytest.mark.parametrize("name,upto", [("walled_rollout_10_w30", None), ("walled_rollout_18_zero", None), ("walled_rollout_13_w33", None),
                                       ("walled_rollout_8_w28", None), ("walled_rollout_9_w29", 1000), ("scene0_zero", 700), ("scene1_w5", 575)])
def test_gpu_equals_the_reference_rollout_directly(name, upto):
    """The CUDA path (reference table order, Jacobi) against the reference's own rollout — no oracle in between —
    wherever the reference's Gauss-Seidel order and BVH false negatives never mattered: every kept frame of the fixture,
    the Welford statistics, the frame count and the comms.result text."""
    gd = np.load(os.path.join(GOLD, name + ".npz"))
    m_cv = gd["m_cv"] if "m_cv" in gd else None
    m_f = gd["m_f"] if "m_f" in gd else None
    pos, goal, md, rng, walls = lcb.scene_full(int(gd["scene"]), int(gd["seed"]))
    assert np.array_equal(pos, gd["pos0"]) and np.array_equal(goal, gd["goal"])
    g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, walls=walls, order="reference")
    at = 0
    for k, f in enumerate(gd["frames"]):
        if upto is not None and f > upto:
            break
        g.step(int(f) - at)
        at = int(f)
        st = g.state()
        for key in KEYS:
            assert np.array_equal(st[key], gd[key][k], equal_nan=True), f"{name} frame {f}: {key} differs from the reference"
        sc = g.scores()
        ra, rs_, av, nd = gd["progress"][k]
        assert (np.float32(sc["run_avg"]), np.float32(sc["run_std"]), sc["n_done"]) == (np.float32(ra), np.float32(rs_), int(nd))
    if upto is None:
        assert int(gd["complete"])
        frames = g.run_to_completion()
        assert frames == int(gd["frames_executed"])
        sc = g.scores()
        got = np.array([sc[k] for k in ("avg_velocity", "group_time", "run_avg", "run_std")], np.float32)
        assert np.array_equal(got, gd["scores"], equal_nan=True) and sc["n_done"] == int(gd["n_done"])
        assert ("%g" % got[0], "%g" % got[1]) == tuple(str(x) for x in gd["result_file"])


@pytest.mark.parametrize("scene,seed", [(9, 0), (18, 2), (13, 1), (3, 0)])
def test_walled_batch_equals_single_crowd(scene, seed):
    """k_batch_rollout with walls (one layout per batch, per-scenario roadmaps cloned at the first replan): 6 particles of a
    walled catalogue scene, rolled out to completion by relaunching while scenarios wait for the host planner, against the
    single-crowd path (itself held against O-canon above): frames, scores, final state, plans, wall hits, replans."""
    pos, goal, md, rng, walls = lcb.scene_full(scene, seed)
    P = 6
    m_cv, m_f = wl.weights(100 + scene, batch=P)
    m_cv[0] = 0
    m_f[0] = 0
    b = lcb.Batch(np.broadcast_to(pos, (P,) + pos.shape).copy(), np.broadcast_to(goal, (P,) + goal.shape).copy(), m_cv, m_f,
                  np.full(P, md, np.float32), np.full(P, rng, np.uint32), walls=walls, order="reference")
    part = b.run(max_frames=150)     # resumable: a bounded first leg
    assert (part["frames"] == 150).all()
    out = b.run()
    bs, bw = b.state(), b.walled()
    replans = 0
    for k in range(P):
        g = lcb.Sim(pos, goal, m_cv[k], m_f[k], max_duration=md, rng_state=rng, walls=walls, order="reference")
        frames = g.run_to_completion()
        sc = g.scores()
        assert out["frames"][k] == frames and out["n_done"][k] == sc["n_done"], (k, out["frames"][k], frames)
        assert out["group_time"][k] == np.float32(sc["group_time"])
        assert out["avg_velocity"][k] == np.float32(sc["avg_velocity"]) or (np.isnan(out["avg_velocity"][k]) and np.isnan(sc["avg_velocity"]))
        _same({key: bs[key][k] for key in KEYS}, g.state(), f"scenario {k}")
        gw = g.walled()
        assert np.array_equal(bw["n_plan"][k], gw["n_plan"]) and bw["wall_hits"][k] == gw["wall_hits"]
        replans += gw["replans"]
    assert bw["replans"] == replans


def test_walled_error_paths():
    pos, goal, md, rng, walls = lcb.scene_full(3, 0)
    with pytest.raises(lcb.LcError):  # a decomposed crowd cannot have walls (replans pause the frame for the host planner)
        d = lcb.api._Desc(len(pos), lcb.api._p(pos), lcb.api._p(goal), 0.15, md, rng, None, None, 0, 0, 1, lcb.api._p(walls), 0, None, None)
        dom = lcb.api._DomainDesc(0, 2, None, 0, 0, None, None)
        import ctypes as C
        h = C.c_void_p()
        lcb.api._ck(lcb.lib().lc_dd_create(C.byref(d), C.byref(dom), C.byref(h)))
    pos, goal, md, rng, walls = lcb.scene_full(6, 0)
    with pytest.raises(lcb.LcError):
        lcb.Sim(pos, goal, walls=walls, wall_order=np.zeros(len(walls), np.int32))  # not a permutation
