"""GPU parity tests proper: the CUDA path (through the C ABI, liblearncomms.so) against O-canon on the same inputs.
Integer outputs (neighbour lists {j : ttc_ij <= 5}, overlap-pair and collision-hit counts, comm partners, done flags)
and — because the kernels use the oracle's operation order with FMA contraction off — all float state are compared
BIT-EXACT (tolerance 0 ulp, stricter than the 4-ulp bound BASELINE.md allows)."""
import numpy as np
import pytest

import _oracle
import learn_comms_b200 as lcb
from learn_comms_b200 import workloads as wl

pytestmark = pytest.mark.gpu

KEYS = ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")


def assert_state_equal(g, o, tag):
    for k in KEYS:
        x, y = g[k], o[k]
        same = np.array_equal(x, y, equal_nan=True) if x.dtype.kind == "f" else np.array_equal(x, y)
        if not same:
            d = np.abs(x.astype(np.float64) - y.astype(np.float64))
            raise AssertionError(f"{tag}: {k} differs, max abs {np.nanmax(d):.3e}, {(d > 0).sum()} elements")


def lockstep(pos, goal, md, rng, m_cv, m_f, frames, comm=True, grid=False, lists=True, check_every=1, stencil=0, features=0):
    g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, comm=comm, debug_lists=lists, stencil=stencil, features=features)
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, comm=comm, grid=grid, features=features)
    o.record_neighbors(lists)
    f = 0
    while f < frames and not o.finished:
        k = min(check_every, frames - f)
        g.step(k)
        o.step(k)
        f += k
        assert_state_equal(g.state(), o.state(), f"frame {f}")
        if comm:
            assert np.array_equal(g.partners(), o.partners()), f"frame {f}: comm partners differ"
        assert np.array_equal(g.neighbor_counts(), o.neighbor_counts()), f"frame {f}: neighbour counts differ"
        if lists:
            gp, gi = g.neighbor_lists()
            op, oi = o.neighbor_lists()
            assert np.array_equal(gp, op) and np.array_equal(gi, oi), f"frame {f}: neighbour lists differ"
        cg = g.counters()
        assert cg["overlap_pairs"] == o.overlap_pairs, f"frame {f}: overlap-pair count"
        assert cg["collision_hits"] == o.collision_hits, f"frame {f}: collision-hit count"
        assert g.rng == o.rng
    return g, o


@pytest.mark.parametrize("scene_id,wseed", [(0, None), (0, 1), (1, 5), (4, 2), (5, 7), (12, 3)])
def test_scene_rollout_bit_exact(scene_id, wseed):
    """Full rollouts of the reference's wall-free scenes to completion: per-frame state, lists and final scores."""
    pos, goal, md, rng = lcb.scene(scene_id, 0)
    m_cv, m_f = (None, None) if wseed is None else wl.weights(wseed)
    g, o = lockstep(pos, goal, md, rng, m_cv, m_f, frames=10 ** 6, check_every=1 if scene_id in (0, 4) else 7)
    assert g.finished and o.finished and g.frame == o.frame
    gs, os_ = g.scores(), o.scores()
    for k in gs:
        assert gs[k] == os_[k] or (np.isnan(gs[k]) and np.isnan(os_[k])), (k, gs, os_)


@pytest.mark.parametrize("features", [lcb.FEAT_NO_PROX | lcb.FEAT_NO_SOFTSIGN, lcb.FEAT_NO_PROX, lcb.FEAT_PLAIN, lcb.FEAT_PLAIN | lcb.FEAT_NO_SOFTSIGN])
def test_feature_set_variants(features):
    """comm::x's #define variants (comm.cpp:22-28): O-canon's flags are pinned against the reference built with PROX and
    SOFTSIGN_GOAL off (tests/golden/scene4_w2_noprox_nosoftsign.npz); the CUDA path follows O-canon bit for bit in each."""
    pos, goal, md, rng = lcb.scene(4, 0)
    m_cv, m_f = wl.weights(2)
    g, o = lockstep(pos, goal, md, rng, m_cv, m_f, frames=10 ** 6, check_every=11, features=features)
    assert g.finished and g.frame == o.frame and g.scores() == o.scores()
    # and the batched kernel
    b = lcb.Batch(pos[None], goal[None], m_cv[None], m_f[None], [md], [rng], features=features)
    out = b.run()
    assert out["frames"][0] == o.frame and out["group_time"][0] == np.float32(o.scores()["group_time"])


def test_gpu_equals_reference_variant_golden_directly():
    """... and against the reference variant's own rollout, no oracle in between (reference table order; scene 8 has no
    agent contacts, so the Jacobi pass equals the reference's)."""
    import os
    gd = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scene8_w28_noprox_nosoftsign.npz"))
    pos, goal, md, rng, walls = lcb.scene_full(8, 0)
    g = lcb.Sim(pos, goal, gd["m_cv"], gd["m_f"], max_duration=md, rng_state=rng, walls=walls, order="reference", features=int(gd["features"]))
    at = 0
    for k, f in enumerate(gd["frames"]):
        g.step(int(f) - at)
        at = int(f)
        st = g.state()
        for key in KEYS:
            assert np.array_equal(st[key], gd[key][k], equal_nan=True), f"frame {f}: {key}"
    sc = g.scores()
    assert ("%g" % np.float32(sc["avg_velocity"]), "%g" % np.float32(sc["group_time"])) == tuple(str(x) for x in gd["result_file"])


@pytest.mark.parametrize("name", ["comms_good", "comms_82", "comms_g", "comms_hand"])
def test_legacy_config_files_run(name):
    """BASELINE.json configs[0]: the reference's legacy "3 5" parameter files (HEAD's own parser aborts on them) through
    lc_read_config_legacy -> plain feature set, scene 0 to completion, CUDA path == O-canon."""
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "legacy_configs", name + ".config")
    m_cv, m_f, feat = lcb.read_config_legacy(path)
    assert feat == lcb.FEAT_PLAIN and m_cv[0].any() and not m_cv[1:].any() and not m_f[:, 1:].any()
    with pytest.raises(lcb.LcError):
        lcb.read_config(path)   # HEAD's format: the file is rejected, as the reference's parser rejects it
    pos, goal, md, rng = lcb.scene(0, 0)
    g, o = lockstep(pos, goal, md, rng, m_cv, m_f, frames=10 ** 6, check_every=50, features=feat, lists=False)
    assert g.finished and g.frame == o.frame and g.scores() == o.scores()
    assert np.isfinite(g.scores()["group_time"])


def test_run_to_completion_matches_stepping():
    pos, goal, md, rng = lcb.scene(0, 0)
    m_cv, m_f = wl.weights(11)
    a = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
    frames = a.run_to_completion()
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
    k = o.step(10 ** 6)
    assert frames == k == o.frame
    assert a.scores() == o.scores()


def test_timeout_penalty_and_state_round_trip():
    """A rollout that ends through the time-out penalty (demo.cpp:691-700: finish times of the not-done agents folded in
    id order — the only in-loop reader of the id -> slot table, which k_reorder maintains only around that frame), with
    a get_state / set_state round trip (rebuilds the table) in the middle: frames, scores and state equal the oracle's."""
    pos, goal, _, rng = wl.plaza(1500, 90.0, seed=33, goal_range=60.0)
    m_cv, m_f = wl.weights(13)
    md = 2.5
    g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
    g.step(70)
    o.step(70)
    st = g.state()
    g.set_state(**st)
    g.step(40)
    o.step(40)
    assert_state_equal(g.state(), o.state(), "after the round trip")
    frames = g.run_to_completion()
    o.step(10 ** 6)
    assert frames == o.frame and 148 <= frames <= 152  # dt * frame first exceeds 2.5 s in float arithmetic
    gs, os_ = g.scores(), o.scores()
    assert gs == os_, (gs, os_)
    assert_state_equal(g.state(), o.state(), "final")


def test_nocomm_flavour():
    """nocomm_norender (main.cpp:130-132): comm::run compiled out."""
    pos, goal, md, rng = lcb.scene(4, 0)
    lockstep(pos, goal, md, rng, None, None, frames=400, comm=False, check_every=5)


def test_circle_1k_grid_vs_bruteforce_oracle():
    """C2 (1 000 agents, r = 80): GPU uniform grid vs the oracle's brute-force all-pairs loop (LMP.cpp:288-304)."""
    pos, goal, md, rng = wl.circle(1000, 80.0, seed=1)
    m_cv, m_f = wl.weights(1, 0.1)
    lockstep(pos, goal, md, rng, m_cv, m_f, frames=60, check_every=6)


@pytest.mark.parametrize("stencil", [1, 2, 3])
def test_stencil_variants_same_result(stencil):
    """Cell side = cutoff / stencil: the result must not depend on the grid resolution."""
    pos, goal, md, rng = wl.plaza(4000, 130.0, seed=9, goal_range=30.0)
    m_cv, m_f = wl.weights(2)
    lockstep(pos, goal, md, rng, m_cv, m_f, frames=40, check_every=8, stencil=stencil)


def test_pair_list_overflow_falls_back_to_full_scan():
    """A dense patch inside a sparse crowd: some agents collect more filter survivors than PAIR_CAP, k_force then scans
    the whole disc for them; everybody else takes the list path.  Same result either way."""
    pos, goal, md, rng = wl.plaza(3000, 110.0, seed=21, goal_range=30.0)
    rs = np.random.RandomState(3)
    pos[:300] = rs.uniform(-2.5, 2.5, (300, 2)).astype(np.float32)
    m_cv, m_f = wl.weights(4)
    lockstep(pos, goal, md, rng, m_cv, m_f, frames=30, check_every=5)


def test_far_apart_clusters_grid_cap():
    """Two clusters 8 km apart (in x and y): cutoff-sized cells would need ~1e7 grid cells, so the grid set-up has to
    widen the cells (first back to square, then doubling) until they fit the 2^20-cell table; same result."""
    pa, ga, _, _ = wl.plaza(120, 30.0, seed=12, goal_range=20.0)
    pb, gb, _, _ = wl.plaza(120, 30.0, seed=13, goal_range=20.0)
    off = np.float32(8000.0)
    pos = np.concatenate([pa, pb + off]).astype(np.float32)
    goal = np.concatenate([ga, gb + off]).astype(np.float32)
    m_cv, m_f = wl.weights(6)
    lockstep(pos, goal, 20.0, 3, m_cv, m_f, frames=40, check_every=8)


def test_dense_crowd_collisions():
    """Agents packed below contact distance: exercises the overlap predicate and push-out (physics.cpp:23-36,
    BVH.cpp:106-120) with many simultaneous hits, and t = 0 / mag = 20 TTC forces."""
    rs = np.random.RandomState(4)
    n = 600
    pos = rs.uniform(-3, 3, (n, 2)).astype(np.float32)
    goal = rs.uniform(-3, 3, (n, 2)).astype(np.float32)
    m_cv, m_f = wl.weights(8)
    g, o = lockstep(pos, goal, 30.0, 1, m_cv, m_f, frames=30, check_every=1)
    assert o.overlap_pairs > 0 or o.collision_hits > 0


def test_teacher_forced_single_step():
    """set_state -> one frame -> compare, from states sampled along an oracle trajectory."""
    pos, goal, md, rng = lcb.scene(12, 0)
    m_cv, m_f = wl.weights(21)
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
    g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
    for hop in (1, 50, 133, 200):
        o.step(hop)
        st = o.state()
        g.set_state(**st)
        g.rng = o.rng
        o2 = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
        o2.set_state(**st)
        o2.rng = o.rng
        g.step(1)
        o2.step(1)
        assert_state_equal(g.state(), o2.state(), f"teacher-forced after {hop}")
        assert np.array_equal(g.partners(), o2.partners())
        assert np.array_equal(g.neighbor_counts(), o2.neighbor_counts())


def test_exchange_state_equals_get_then_set():
    """lc_exchange_state == lc_get_state followed by lc_set_state (the two copies overlap; the result may not differ):
    teacher forcing along an oracle trajectory, reference table order, reading every frame's result back."""
    pos, goal, md, rng = lcb.scene(5, 0)
    m_cv, m_f = wl.weights(23)
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
    a = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, order="reference")
    b = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, order="reference")
    out = {k: np.zeros_like(v) for k, v in a.state().items()}
    for hop in (3, 40, 41, 90):
        o.step(hop)
        forced = o.state()
        a.step(1)
        b.step(1)
        want = b.state()
        b.set_state(**forced)
        got = a.exchange_state(out, **forced)
        assert_state_equal(got, want, f"exchange_state out tables at hop {hop}")
        assert_state_equal(a.state(), b.state(), f"state after exchange_state at hop {hop}")
        a.rng = b.rng = o.rng
    a.step(5)
    b.step(5)
    assert_state_equal(a.state(), b.state(), "five frames on")
    part = {"pos": out["pos"], "vel": out["vel"]}   # partial tables: the others keep what they were
    a.exchange_state(part, pos=forced["pos"])
    b_st = b.state()
    assert np.array_equal(part["pos"], b_st["pos"]) and np.array_equal(part["vel"], b_st["vel"])
    b.set_state(pos=forced["pos"])
    assert_state_equal(a.state(), b.state(), "partial exchange")


def test_corridor_grid_oracle_medium():
    """C4 shape at 1/10 scale (10 000 agents) against the grid-accelerated oracle (itself checked against brute
    force in the CPU suite)."""
    pos, goal, md, rng = wl.corridor(cols=100, rows=50, gap=6.0, seed=4)
    m_cv, m_f = wl.weights(4, 0.5)
    lockstep(pos, goal, md, rng, m_cv, m_f, frames=24, grid=True, lists=True, check_every=8)


def test_full_size_properties_plaza():
    """C5 full size (1M agents): size-independent properties — determinism (two handles, same bits), permutation
    invariance of the per-agent result w.r.t. slot order is implied by determinism after re-sorting, id bijection,
    speed clamp |vel_forehalf| <= 1 up to push-out, and agreement with the grid oracle on a 3-frame prefix."""
    pos, goal, md, rng = wl.plaza(1_000_000, 2000.0, seed=5)
    m_cv, m_f = wl.weights(5, 0.5)
    a = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, free_run=True)
    b = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, free_run=True, stencil=2)
    a.step(3)
    b.step(3)
    sa, sb = a.state(), b.state()
    assert_state_equal(sa, sb, "stencil 1 vs 2 at 1M")
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, grid=True)
    o.step(3, fixed=True)
    assert_state_equal(sa, o.state(), "1M agents vs grid oracle, 3 frames")
    assert np.array_equal(a.partners(), o.partners())
    assert np.array_equal(a.neighbor_counts(), o.neighbor_counts())
    a.step(60)
    s = a.state()
    assert np.isfinite(s["pos"]).all()
    assert (np.hypot(s["vfh"][:, 0], s["vfh"][:, 1]) <= 1.0 + 1e-3).all()


@pytest.mark.parametrize("n", [33, 201])
def test_odd_agent_count_state_tables(n):
    """Odd agent counts: every sub-table of the lc_get_state / lc_set_state staging must stay aligned for its vector
    type (float2 tables at offsets that are odd multiples of 4 bytes fault with cudaErrorMisalignedAddress)."""
    pos, goal, md, rng = wl.plaza(n, 14.0, seed=n, goal_range=10.0)
    m_cv, m_f = wl.weights(n)
    g, o = lockstep(pos, goal, 30.0, rng, m_cv, m_f, frames=20, check_every=5)
    st = g.state()
    g.set_state(**st)
    g.step(5)
    o.step(5)
    assert_state_equal(g.state(), o.state(), f"n = {n} after a set_state round trip")


def test_device_error_surfaces_without_run_to_completion():
    """A host that only steps and reads tables (the loop INTEGRATION.md recommends) must see Ctl::error: a non-finite
    position poisons the grid set-up (LC_ERR_STATE) and lc_finished / lc_get_state report it."""
    pos, goal, md, rng = lcb.scene(0, 0)
    g = lcb.Sim(pos, goal, None, None, max_duration=md, rng_state=rng)
    g.step(2)
    st = g.state()
    st["pos"][3] = np.nan
    g.set_state(**st)
    g.step(2)
    with pytest.raises(lcb.LcError) as e:
        _ = g.finished
    assert e.value.code == -3
    with pytest.raises(lcb.LcError):
        g.state()


def test_error_paths():
    pos, goal, md, rng = lcb.scene(0, 0)
    with pytest.raises(lcb.LcError):
        lcb.Sim(pos[:1], goal[:1])  # n == 1: reference UB (comm.cpp:273-285) -> LC_ERR_ARG
    with pytest.raises(lcb.LcError):
        lcb.Sim(pos, goal, device=99)
