"""Host-side global planner of the walled scenes (lc_planner_*, learn_comms_b200/csrc/planner.cpp) against
the reference's own Cspace2d + PRM + GMP::plan_one (oracle/_ref/oref_plan; fixtures by tools/make_golden_planner.py).
The planner runs on the host in the reference too (ai::init, ai.cpp:41-127), so these tests need no GPU.  Bar: roadmap
size, every waypoint list (equal-cost ties included) and min_time identical, bit for bit."""
import importlib.util
import os

import numpy as np
import pytest

import ctypes as C
import subprocess

import _oracle
import _oref
import _walled
import learn_comms_b200 as lcb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "planner.npz")


def _check(walls, start, goal, roadmap, plans):
    p = lcb.Planner(walls, 0.15)
    assert p.roadmap_size() == tuple(int(v) for v in roadmap)
    for i in range(len(start)):  # same order as the reference: the roadmap is stateful (ids, hash-table history)
        wp, mt = p.plan(start[i], goal[i])
        ref_mt, ref_wp = plans[i]
        assert wp.shape == ref_wp.shape and np.array_equal(wp, ref_wp), "agent %d: waypoints differ" % i
        assert np.float32(mt) == np.float32(ref_mt), "agent %d: min_time differs" % i
    p.close()


def test_planner_matches_reference_golden():
    g = np.load(GOLD)
    multi = 0
    for s in g["seeds"]:
        cnt = g[f"cnt{s}"]
        plans = [(g[f"min_time{s}"][i], g[f"wp{s}"][i, :cnt[i]]) for i in range(len(cnt))]
        _check(g[f"walls{s}"], g[f"start{s}"], g[f"goal{s}"], g[f"roadmap{s}"], plans)
        multi += int((cnt >= 3).sum())
    assert multi >= 30  # the fixture holds real detours, not only straight lines


@pytest.mark.skipif(not _oref.plan_available(), reason="oracle/_ref/oref_plan not built (needs /root/reference)")
def test_live_reference_planner_matches():
    spec = importlib.util.spec_from_file_location("mgp", os.path.join(ROOT, "tools", "make_golden_planner.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for seed in (100, 101, 102, 103, 104, 105):
        walls, start, goal = mod.layout(seed, n_agents=60)
        nv, ne, plans = _oref.run_plan(walls, 0.15, start, goal)
        _check(walls, start, goal, (nv, ne), plans)


def test_planner_known_answers():
    """One 4 m block at the origin, agent radius 0.15: inflated half-side 2.15, roadmap points 0.3 m further out."""
    p = lcb.Planner([[0, 0, 4, 4]], 0.15)
    assert p.roadmap_size() == (8, 24)  # each point sees its two neighbours on the ring, plus the far corner of its face
    wp, mt = p.plan((-8, 3), (8, 3))    # clear of the block: straight line
    assert wp.tolist() == [[8.0, 3.0]] and abs(mt - 15.0) < 1e-6
    wp, mt = p.plan((-8, 0), (8, 0))    # around a corner pair
    assert len(wp) >= 3 and np.allclose(np.abs(wp[:-1, 1]), 2.45, atol=1e-6) and wp[-1].tolist() == [8.0, 0.0]
    expected = 2 * np.hypot(8 - 2.45, 2.45) + 4.9 - 1
    assert abs(mt - expected) < 1e-4
    wp, mt = p.plan((0, 0), (8, 0))     # start inside the wall: unreachable, empty plan, min_time = -1 (ai.cpp:121)
    assert len(wp) == 0 and mt == -1.0
    wp, mt = p.plan((-8, 3), (8, 3))    # the temporary start / goal vertices left the roadmap again
    assert p.roadmap_size() == (8, 24) and wp.tolist() == [[8.0, 3.0]]
    # no walls: the plan is the goal (what the wall-free product path assumes)
    q = lcb.Planner(np.zeros((0, 4), np.float32), 0.15)
    wp, mt = q.plan((1, 1), (4, 5))
    assert wp.tolist() == [[4.0, 5.0]] and abs(mt - 4.0) < 1e-6


def test_planner_clone_is_independent_and_resolves_ties_identically():
    """lc_planner_clone: the walled batch gives every scenario its own copy of the layout's roadmap (the roadmap is
    stateful: temporary vertices come and go, hash containers remember their history).  A clone must answer a sequence of
    queries exactly like the original would have, and must not see what happens to the original afterwards."""
    rs = np.random.RandomState(3)
    walls = [[x, y, 2.0, 2.0] for x in (-6, -2, 2, 6) for y in (-4, 0, 4)]
    a = lcb.Planner(walls, 0.15)
    warm = [(rs.uniform(-9, 9, 2), rs.uniform(-9, 9, 2)) for _ in range(25)]
    for s_, g_ in warm:
        a.plan(tuple(s_), tuple(g_))
    b = a.clone()
    assert b.roadmap_size() == a.roadmap_size()
    queries = [(rs.uniform(-9, 9, 2), rs.uniform(-9, 9, 2)) for _ in range(40)]
    ra = [a.plan(tuple(s_), tuple(g_)) for s_, g_ in queries]
    for _ in range(10):   # the original moves on; the clone is asked the same 40 questions afterwards
        a.plan((-9.0, -9.0), (9.0, 9.0))
    rb = [b.plan(tuple(s_), tuple(g_)) for s_, g_ in queries]
    for (wa, ta), (wb, tb) in zip(ra, rb):
        assert np.array_equal(wa, wb) and np.float32(ta) == np.float32(tb)
    fresh = lcb.Planner(walls, 0.15)   # ... and like a planner that replays the whole history
    for s_, g_ in warm:
        fresh.plan(tuple(s_), tuple(g_))
    for (s_, g_), (wa, ta) in zip(queries, ra):
        wf, tf = fresh.plan(tuple(s_), tuple(g_))
        assert np.array_equal(wa, wf) and np.float32(ta) == np.float32(tf)
    b.close()
    assert a.roadmap_size()[0] > 0   # closing a clone leaves the original intact


def test_planner_error_paths():
    with pytest.raises(lcb.LcError):
        lcb.Planner([[0, 0, 1, 1]], agent_radius=0.0)
    p = lcb.Planner([[0, 0, 4, 4]], 0.15)
    with pytest.raises(lcb.LcError):
        p.plan((-8, 0), (8, 0), cap=1)  # LC_ERR_CAPACITY


# ---------------------------------------------------------------- LMP::lookahead over the planned waypoint lists
def _host_build_of_device_header(tmp_path):
    so = tmp_path / "walls_host.so"
    subprocess.check_call(["g++", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math", "-w", "-x", "c++",
                           os.path.join(ROOT, "tests", "walls_host_shim.cpp"), "-o", str(so)])
    return C.CDLL(str(so))


def _device_header_lookahead(L, walls, recs):
    walls = np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
    recs = np.ascontiguousarray(recs, _oref.LOOK_REC)
    out = np.zeros(len(recs), _oref.LOOK_OUT)
    L.walls_host_lookahead(C.c_int32(len(walls)), walls.ctypes.data_as(C.c_void_p), C.c_float(0.15), C.c_int32(len(recs)),
                           recs.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    return out


def test_lookahead_matches_reference_golden(tmp_path):
    """LMP::lookahead (LMP.cpp:84-113) with multi-waypoint plans and walls: O-canon's restatement and
    the product's device header (walls.cuh lcw::lookahead, compiled for the host) both return the reference's num_done
    and local_goal bit for bit on 2 454 records (positions scattered around the planned paths, every num_done)."""
    L = _host_build_of_device_header(tmp_path)
    g = np.load(GOLD)
    total = advanced = 0
    for s in g["seeds"]:
        rin, rout = g[f"look_in{s}"], g[f"look_out{s}"]
        assert _oracle.lookahead(g[f"walls{s}"], 0.15, rin).tobytes() == rout.tobytes()
        assert _device_header_lookahead(L, g[f"walls{s}"], rin).tobytes() == rout.tobytes()
        total += len(rin)
        advanced += int((rout["num_done"] > rin["num_done"]).sum())
    assert total > 2000 and advanced > 200


@pytest.mark.skipif(not (_oref.look_available() and _oref.plan_available()), reason="oracle/_ref harnesses not built")
def test_live_reference_lookahead_matches(tmp_path):
    spec = importlib.util.spec_from_file_location("mgp", os.path.join(ROOT, "tools", "make_golden_planner.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    L = _host_build_of_device_header(tmp_path)
    for seed in (200, 201, 202):
        walls, start, goal = mod.layout(seed, n_agents=60)
        _, _, plans = _oref.run_plan(walls, 0.15, start, goal)
        recs = mod.look_records(seed, start, plans, per_plan=10)
        ref = _oref.run_look(walls, 0.15, recs)
        assert _oracle.lookahead(walls, 0.15, recs).tobytes() == ref.tobytes()
        assert _device_header_lookahead(L, walls, recs).tobytes() == ref.tobytes()


# ---------------------------------------------------------------- the reference's own walled rollout (scene 3)
def _golden_rollout():
    g = np.load(os.path.join(ROOT, "tests", "golden", "walled_scene3.npz"))
    frames = [dict(frame=int(g["frame"][k]), pos=g["pos"][k], goal_dist=g["goal_dist"][k],
                   num_done=g["num_done"][k].astype(np.int32), local_goal=g["local_goal"][k]) for k in range(len(g["frame"]))]
    return dict(n=int(g["n"]), order=g["order"], pos0=g["pos0"], goal=g["goal"], min_time=g["min_time"], frames=frames)


def test_walled_scene3_initial_plans_match_reference():
    """ai::init on the reference's scene 3 (60 agents around one 4 m pillar): planning the agents in the reference's
    table order gives every agent's min_time bit for bit (ai.cpp:107-127)."""
    r = _golden_rollout()
    p = lcb.Planner(_walled.SCENE3_WALLS, 0.15)
    mt = np.zeros(r["n"], np.float32)
    longest = 0
    for i in r["order"]:
        wp, mt[i] = p.plan(r["pos0"][i], r["goal"][i])
        longest = max(longest, len(wp))
    assert np.array_equal(mt, r["min_time"]) and longest == 3


def test_walled_scene3_waypoint_bookkeeping_matches_reference_rollout():
    """700 consecutive frames of the reference's own scene-3 rollout, teacher-forced on its positions: sight-line gate,
    lost_frames replans (from frame 361), the two-per-frame replan limit in table order, an unreachable replan (agent
    inside the inflated pillar: no plan, local_goal = own position) and LMP::lookahead give the reference's num_done
    and local_goal for all 60 agents in every frame."""
    first_bad, replans = _walled.follow(_golden_rollout(), _walled.SCENE3_WALLS)
    assert first_bad is None, "differs from the reference at frame %s" % first_bad
    assert replans >= 20


@pytest.mark.skipif(not _oref.available(), reason="oracle/_ref/oref not built (needs /root/reference)")
def test_live_walled_scene3_full_rollout_bookkeeping():
    """The same replay over a complete live rollout (to all_done) with other weights."""
    from learn_comms_b200 import workloads as wl
    m_cv, m_f = wl.weights(9)
    r = _oref.run(3, 0, m_cv, m_f, dump_every=1)
    first_bad, replans = _walled.follow(r, _walled.SCENE3_WALLS)
    assert first_bad is None, "differs from the reference at frame %s" % first_bad
    assert r["n_walls"] == 1 and replans > 0
