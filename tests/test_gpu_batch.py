"""GPU parity of the batched-rollout kernel (lc_batch_*, one thread block per scenario) against O-canon, brute-force
mode, one oracle simulation per scenario.  Same bar as the single-crowd tests: every float of the state, the comm
partners, neighbour counts, overlap / collision-hit totals, frame counts and final scores BIT-EXACT."""
import numpy as np
import pytest

import _oracle
import learn_comms_b200 as lcb
from learn_comms_b200 import workloads as wl

pytestmark = pytest.mark.gpu

KEYS = ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")


def _eq(x, y):
    return np.array_equal(x, y, equal_nan=True) if x.dtype.kind == "f" else np.array_equal(x, y)


def _oracles(pos, goal, m_cv, m_f, md, rng, comm=True):
    return [_oracle.CanonSim(pos[s], goal[s], m_cv[s], m_f[s], max_duration=float(md[s]), rng_state=int(rng[s]), comm=comm)
            for s in range(pos.shape[0])]


def _score_eq(a, b):
    return a == b or (np.isnan(a) and np.isnan(b))


def test_batch_lockstep_bit_exact():
    """16 scenarios (8 layouts x 2 particles), stepped 25 frames at a time for 500 frames: state + diagnostics."""
    pos, goal, m_cv, m_f, md, rng = wl.pso_batch(16, 64, seed=11)
    b = lcb.Batch(pos, goal, m_cv, m_f, md, rng)
    os_ = _oracles(pos, goal, m_cv, m_f, md, rng)
    overlap = np.zeros(16, np.int64)
    hits = np.zeros(16, np.int64)
    for chunk in range(20):
        out = b.run(25)
        st, dg = b.state(), b.diag()
        for s, o in enumerate(os_):
            for _ in range(25):
                if o.finished:
                    break
                o.step(1)
                overlap[s] += o.overlap_pairs
                hits[s] += o.collision_hits
            ost = o.state()
            for k in KEYS:
                assert _eq(st[k][s], ost[k]), f"chunk {chunk} scenario {s}: {k} differs"
            assert out["frames"][s] == o.frame
            assert np.array_equal(dg["partner"][s], o.partners()), f"chunk {chunk} scenario {s}: partners"
            assert np.array_equal(dg["neighbor_count"][s], o.neighbor_counts()), f"chunk {chunk} scenario {s}: neighbour counts"
        assert np.array_equal(dg["overlap_pairs"], overlap) and np.array_equal(dg["collision_hits"], hits)


@pytest.mark.parametrize("comm", [True, False])
def test_batch_full_rollouts_scores(comm):
    """64 scenarios to completion in one launch: frames, scores and final state equal the oracle's, per scenario."""
    pos, goal, m_cv, m_f, md, rng = wl.pso_batch(64, 64, seed=5)
    md[::7] = 8.0  # some scenarios end through the timeout penalty (demo.cpp:691-700)
    b = lcb.Batch(pos, goal, m_cv, m_f, md, rng, comm=comm)
    out = b.run()
    assert b.n_finished == 64 and b.launches == 1
    st = b.state()
    for s, o in enumerate(_oracles(pos, goal, m_cv, m_f, md, rng, comm=comm)):
        o.step(10 ** 6)
        sc = o.scores()
        assert out["frames"][s] == o.frame, (s, out["frames"][s], o.frame)
        assert _score_eq(out["avg_velocity"][s], np.float32(sc["avg_velocity"])), (s, out["avg_velocity"][s], sc)
        assert _score_eq(out["group_time"][s], np.float32(sc["group_time"])), (s, out["group_time"][s], sc)
        assert out["n_done"][s] == sc["n_done"]
        ost = o.state()
        for k in KEYS:
            assert _eq(st[k][s], ost[k]), f"scenario {s}: final {k} differs"


def test_batch_matches_single_crowd_path_on_reference_scenes():
    """The reference's own scenes 0 / 4 / 5 (60 / 56 / 56 agents: not a multiple of the warp size) through the batch kernel
    == the single-crowd kernels == the oracle."""
    for scene_id in (0, 4, 5):
        pos, goal, md, rng = lcb.scene(scene_id, 0)
        m_cv, m_f = wl.weights(scene_id + 1)
        b = lcb.Batch(pos[None], goal[None], m_cv[None], m_f[None], [md], [rng])
        out = b.run()
        g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng)
        frames = g.run_to_completion()
        gs = g.scores()
        assert out["frames"][0] == frames
        assert _score_eq(out["group_time"][0], np.float32(gs["group_time"])) and _score_eq(out["avg_velocity"][0], np.float32(gs["avg_velocity"]))
        st, gst = b.state(), g.state()
        for k in KEYS:
            assert _eq(st[k][0], gst[k]), f"scene {scene_id}: {k}"


def test_batch_dense_collisions_and_odd_sizes():
    """Packed random crowds (initial overlaps -> collision pass, near-list overflow) at agent counts 2, 33, 200, 256."""
    for n, side in ((2, 0.2), (33, 1.2), (200, 3.0), (256, 3.5)):
        rs = np.random.RandomState(n)
        S = 3
        pos = rs.uniform(-side, side, (S, n, 2)).astype(np.float32)
        goal = rs.uniform(-10, 10, (S, n, 2)).astype(np.float32)
        m_cv, m_f = wl.weights(n, 1.0, batch=S)
        md, rng = np.full(S, 3.0, np.float32), np.arange(1, S + 1).astype(np.uint32)
        b = lcb.Batch(pos, goal, m_cv, m_f, md, rng)
        os_ = _oracles(pos, goal, m_cv, m_f, md, rng)
        hits = np.zeros(S, np.int64)
        for chunk in range(6):
            b.run(10)
            st, dg = b.state(), b.diag()
            for s, o in enumerate(os_):
                for _ in range(10):
                    o.step(1)
                    hits[s] += o.collision_hits
                ost = o.state()
                for k in KEYS:
                    assert _eq(st[k][s], ost[k]), f"n={n} chunk {chunk} scenario {s}: {k} differs"
            assert np.array_equal(dg["collision_hits"], hits), (n, dg["collision_hits"], hits)
        assert hits.sum() > 0 or n == 2


def test_batch_reset_reproduces_and_reweights():
    pos, goal, m_cv, m_f, md, rng = wl.pso_batch(32, 64, seed=2)
    b = lcb.Batch(pos, goal, m_cv, m_f, md, rng)
    a = b.run()
    b.reset()
    c = b.run()
    for k in a:
        assert _eq(a[k], c[k]), k
    m_cv2, m_f2 = wl.weights(99, 1.0, batch=32)
    b.reset(m_cv2, m_f2)
    d = b.run()
    e = lcb.Batch(pos, goal, m_cv2, m_f2, md, rng).run()
    for k in d:
        assert _eq(d[k], e[k]), k
    assert not _eq(a["group_time"], d["group_time"])


def test_batch_error_paths():
    pos, goal, m_cv, m_f, md, rng = wl.pso_batch(2, 64)
    with pytest.raises(lcb.LcError) as e:
        lcb.Batch(pos[:, :1], goal[:, :1], m_cv, m_f, md, rng)
    assert e.value.code == -1
    big = np.zeros((1, 300, 2), np.float32)
    with pytest.raises(lcb.LcError):
        lcb.Batch(big, big, m_cv[:1], m_f[:1], md[:1], rng[:1])
    # coincident agents: nobody has a comm partner (reference UB, comm.cpp:273-285) -> LC_ERR_STATE, not a crash
    z = np.zeros((1, 4, 2), np.float32)
    b = lcb.Batch(z, z + 5, m_cv[:1], m_f[:1], md[:1], rng[:1])
    with pytest.raises(lcb.LcError) as e:
        b.run(1)
    assert e.value.code == -3


def test_batch_full_size_properties():
    """BASELINE.json configs[2] at full size (4 096 scenarios x 64 agents): properties that need no oracle run —
    determinism, invariance under sharding (what the multi-GPU path relies on), every scenario finishes within the
    reference's 45 s limit (+1 frame), finite scores, and a 32-scenario sample against the oracle."""
    pos, goal, m_cv, m_f, md, rng = wl.pso_batch(4096, 64, seed=3)
    full = lcb.Batch(pos, goal, m_cv, m_f, md, rng)
    a = full.run()
    assert full.n_finished == 4096
    full.reset()
    b = full.run()
    for k in a:
        assert _eq(a[k], b[k]), f"not deterministic: {k}"
    lo = lcb.Batch(pos[:1500], goal[:1500], m_cv[:1500], m_f[:1500], md[:1500], rng[:1500]).run()
    hi = lcb.Batch(pos[1500:], goal[1500:], m_cv[1500:], m_f[1500:], md[1500:], rng[1500:]).run()
    for k in a:
        assert _eq(a[k], np.concatenate([lo[k], hi[k]])), f"sharding changes {k}"
    assert a["frames"].max() <= 2701 and a["frames"].min() > 60
    assert np.isfinite(a["group_time"]).all() and (a["n_done"] >= 64).all()
    idx = np.arange(0, 4096, 128)
    for s in idx:
        o = _oracle.CanonSim(pos[s], goal[s], m_cv[s], m_f[s], max_duration=float(md[s]), rng_state=int(rng[s]))
        o.step(10 ** 6)
        sc = o.scores()
        assert a["frames"][s] == o.frame and _score_eq(a["group_time"][s], np.float32(sc["group_time"])) \
            and _score_eq(a["avg_velocity"][s], np.float32(sc["avg_velocity"])), s
