"""Spatial domain decomposition (lc_dd_*): a crowd split into x strips, one rank per strip, boundary agents exchanged by
the kernels through peer memory every frame.  The bar: the merged state of the ranks equals the single-GPU path
(lc_create) BIT FOR BIT every checked frame, and so do frame counts and final scores.  Ranks live in this process; they
sit on different GPUs when the box has several and share GPU 0 otherwise (the exchange protocol is the same)."""
import numpy as np
import pytest

import learn_comms_b200 as lcb
from learn_comms_b200 import workloads as wl

pytestmark = pytest.mark.gpu

KEYS = ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")


def _devices(world):
    import torch
    k = torch.cuda.device_count()
    return [r % k for r in range(world)]


def _make(pos, goal, md, rng, m_cv, m_f, world, **kw):
    devs = _devices(world)
    sims = [lcb.DomainSim(pos, goal, r, world, m_cv, m_f, max_duration=md, rng_state=rng, device=devs[r], **kw) for r in range(world)]
    lcb.DomainSim.connect_local(sims)
    return sims


def _assert_same(sims, ref, tag):
    a, b = lcb.DomainSim.merged_state(sims), ref.state()
    for k in KEYS:
        same = np.array_equal(a[k], b[k], equal_nan=True) if a[k].dtype.kind == "f" else np.array_equal(a[k], b[k])
        if not same:
            bad = np.flatnonzero((a[k] != b[k]).reshape(a[k].shape[0], -1).any(1))
            raise AssertionError(f"{tag}: {k} differs for {bad.size} agents, first ids {bad[:8]}")


def _crowd(n, w, h, seed, goal_range=None):
    rs = np.random.RandomState(seed)
    m = int(np.ceil(np.sqrt(n * w / h)))
    k = int(np.ceil(n / m))
    gx, gy = np.meshgrid(np.arange(m), np.arange(k))
    pts = np.stack([gx.ravel() * (w / m) - w / 2, gy.ravel() * (h / k) - h / 2], 1)[:n]
    pos = (pts + rs.uniform(-0.1, 0.1, (n, 2))).astype(np.float32)
    goal = rs.uniform([-w / 2, -h / 2], [w / 2, h / 2], (n, 2)).astype(np.float32) if goal_range is None else \
        (pos + rs.uniform(-goal_range, goal_range, (n, 2))).astype(np.float32)
    return pos, goal


@pytest.mark.parametrize("world,every", [(2, 8), (3, 8), (2, 1), (3, 3), (2, 16)])
def test_domain_lockstep_bit_exact(world, every, monkeypatch):
    """4 000 agents in a 120 x 40 m field, goals all over it (agents cross strip boundaries all the time).  `every` = the
    re-sort period (env LC_REBIN_EVERY): in between, owners, ghosts and slots are frozen and the neighbours only refresh each
    other's ghost records (domain.cuh dd_push_body); 1 = the frame-by-frame routing of every boundary agent."""
    monkeypatch.setenv("LC_REBIN_EVERY", str(every))
    pos, goal = _crowd(4000, 120.0, 40.0, seed=world)
    m_cv, m_f = wl.weights(7, 0.5)
    sims = _make(pos, goal, 200.0, 12345, m_cv, m_f, world)
    monkeypatch.delenv("LC_REBIN_EVERY")   # the single-GPU reference keeps its own default
    ref = lcb.Sim(pos, goal, m_cv, m_f, max_duration=200.0, rng_state=12345)
    migrated = 0
    own0 = np.stack([s.owned() for s in sims])
    for chunk in range(30):
        lcb.DomainSim.step_all(sims, 10)
        ref.step(10)
        _assert_same(sims, ref, f"world {world} frame {10 * (chunk + 1)}")
    own1 = np.stack([s.owned() for s in sims])
    migrated = int((own0 != own1).any(0).sum())
    assert migrated > 0, "no agent changed owner: the test does not exercise migration"
    for s in sims:   # 300 frames: the first one always sorts, then one in `every` (more only if somebody moved too far)
        assert s.resorts() == (300 if every == 1 else 2 + 298 // every), (every, s.resorts())   # (frames 1 and 2 always sort)
    full = ref.state()
    seen = np.zeros(pos.shape[0], np.int32)
    for s in sims:
        n_loc, cap = s.counts()
        assert 0 < n_loc <= cap
        sh, k = s.shard()   # the compact host-facing shard: ids, pos, vel of the rank's own agents
        ids = sh["ids"][:k]
        seen[ids] += 1
        assert np.array_equal(sh["pos"][:k], full["pos"][ids]) and np.array_equal(sh["vel"][:k], full["vel"][ids])
    assert (seen == 1).all()
    # diagnostics of the owners
    part = np.full(pos.shape[0], -2, np.int32)
    cnt = np.full(pos.shape[0], -2, np.int32)
    for s in sims:
        o = s.owned().astype(bool)
        part[o] = s.partners()[o]
        cnt[o] = s.neighbor_counts()[o]
    assert np.array_equal(part, ref.partners()) and np.array_equal(cnt, ref.neighbor_counts())


def test_domain_full_rollout_scores():
    """Scene-12-like random crowd to completion on 2 ranks: frame count, scores (Welford in id order on the scoring rank,
    time-out penalty included) and final state equal the single-GPU path."""
    pos, goal = _crowd(600, 60.0, 30.0, seed=9, goal_range=25.0)
    m_cv, m_f = wl.weights(3, 1.0)
    for md in (60.0, 6.0):  # everybody arrives / the time-out fires first
        sims = _make(pos, goal, md, 777, m_cv, m_f, 2)
        ref = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=777)
        frames = ref.run_to_completion()
        done = False
        while not done:
            lcb.DomainSim.step_all(sims, 64)
            done = all(s.finished for s in sims)
        assert all(s.frame == frames for s in sims), ([s.frame for s in sims], frames)
        _assert_same(sims, ref, f"final state md={md}")
        a, b = lcb.DomainSim.merged_scores(sims), ref.scores()
        for k in b:
            assert a[k] == b[k] or (np.isnan(a[k]) and np.isnan(b[k])), (md, k, a, b)


def test_domain_nocomm_and_dense():
    """Packed crowd (collision pass across the strip boundary) without the comm system."""
    rs = np.random.RandomState(4)   # 0.25 m mean spacing < 2 r: hundreds of initial overlaps on both sides of the cut
    pos = rs.uniform([-12, -4], [12, 4], (3000, 2)).astype(np.float32)
    goal = (pos + rs.uniform(-30, 30, (3000, 2))).astype(np.float32)
    sims = _make(pos, goal, 100.0, 5, None, None, 2, comm=False)   # (push-outs of decimetres: re-sorts every frame at first)
    ref = lcb.Sim(pos, goal, None, None, max_duration=100.0, rng_state=5, comm=False)
    hits = 0
    for chunk in range(12):
        lcb.DomainSim.step_all(sims, 5)
        ref.step(5)
        _assert_same(sims, ref, f"frame {5 * (chunk + 1)}")
        hits += ref.counters()["collision_hits"]
        assert sum(s.counters()["collision_hits"] for s in sims) == ref.counters()["collision_hits"]
        assert sum(s.counters()["overlap_pairs"] for s in sims) >= 0
    assert hits > 0


def test_domain_odd_capacity_shard_tables():
    """Odd slot capacity (the default for 1M agents on 3 ranks is odd): the float2 tables of the lc_dd_get_shard staging
    must stay 8-byte aligned."""
    pos, goal = _crowd(1001, 80.0, 20.0, seed=5)
    m_cv, m_f = wl.weights(2, 0.5)
    sims = _make(pos, goal, 100.0, 9, m_cv, m_f, 2, capacity=1501, halo_capacity=333)
    ref = lcb.Sim(pos, goal, m_cv, m_f, max_duration=100.0, rng_state=9)
    lcb.DomainSim.step_all(sims, 12)
    ref.step(12)
    _assert_same(sims, ref, "odd capacity")
    full = ref.state()
    for s in sims:
        assert s.counts()[1] == 1501
        sh, k = s.shard()
        ids = sh["ids"][:k]
        assert np.array_equal(sh["pos"][:k], full["pos"][ids]) and np.array_equal(sh["vel"][:k], full["vel"][ids])


def test_domain_exchange_capacities_fail_cleanly():
    """The two fixed capacities of the all-rank exchange (crowd.cuh DD_EV_CAP / DD_CAND_CAP = 16 384 per rank and frame):
    (a) a mass arrival — 40 000 agents that are within a metre of their goals from the start, so each rank has 20 000
    done-events to forward in one frame; (b) a crowd that is nothing but outline — 36 000 agents on a circle, every one a
    hull vertex and therefore a comm-partner candidate.  Both must surface LC_ERR_CAPACITY (-5) on the ranks through the
    calls a stepping host makes (lc_finished / lc_get_state), not hang and not go unnoticed.  Below the capacities the same
    shapes run bit-identically to one GPU."""
    rs = np.random.RandomState(8)
    # (a) mass arrival
    n = 40000
    side = 200.0
    m = int(np.ceil(np.sqrt(n)))
    gx, gy = np.meshgrid(np.arange(m), np.arange(m))
    pos = (np.stack([gx.ravel(), gy.ravel()], 1)[:n] * (side / m) - side / 2).astype(np.float32)
    goal = (pos + rs.uniform(-0.3, 0.3, (n, 2))).astype(np.float32)
    sims = _make(pos, goal, 50.0, 3, None, None, 2)
    lcb.DomainSim.step_all(sims, 3)
    codes = []
    for s in sims:
        with pytest.raises(lcb.LcError) as e:
            _ = s.finished
        codes.append(e.value.code)
    assert codes == [-5, -5], codes
    # the same shape below the capacity: 12 000 agents, 6 000 events per rank in one frame
    k = 12000
    sims = _make(pos[:k], goal[:k], 50.0, 3, None, None, 2)
    ref = lcb.Sim(pos[:k], goal[:k], None, None, max_duration=50.0, rng_state=3)
    lcb.DomainSim.step_all(sims, 4)
    ref.step(4)
    _assert_same(sims, ref, "mass arrival below the capacity")
    a, b = lcb.DomainSim.merged_scores(sims), ref.scores()
    assert a["n_done"] == b["n_done"] == k and a["run_avg"] == b["run_avg"] and a["run_std"] == b["run_std"]
    # (b) all outline
    n = 36000
    th = np.arange(n) * (2 * np.pi / n)
    pos = np.stack([2000.0 * np.cos(th), 2000.0 * np.sin(th)], 1).astype(np.float32)
    sims = _make(pos, (-pos).astype(np.float32), 50.0, 3, *wl.weights(4, 0.5), 2)
    lcb.DomainSim.step_all(sims, 2)
    codes = []
    for s in sims:
        with pytest.raises(lcb.LcError) as e:
            s.state()
        codes.append(e.value.code)
    assert codes == [-5, -5], codes
    k = 9000
    th = np.arange(k) * (2 * np.pi / k)
    pos = np.stack([600.0 * np.cos(th), 600.0 * np.sin(th)], 1).astype(np.float32)
    m_cv, m_f = wl.weights(4, 0.5)
    sims = _make(pos, (-pos).astype(np.float32), 50.0, 3, m_cv, m_f, 2)
    ref = lcb.Sim(pos, (-pos).astype(np.float32), m_cv, m_f, max_duration=50.0, rng_state=3)
    lcb.DomainSim.step_all(sims, 6)
    ref.step(6)
    _assert_same(sims, ref, "all-outline crowd below the capacity")


def test_domain_error_paths():
    pos, goal = _crowd(500, 30.0, 30.0, seed=1)
    with pytest.raises(lcb.LcError):
        lcb.DomainSim(pos, goal, 0, 1)
    with pytest.raises(lcb.LcError):  # interior strip narrower than the ghost band
        lcb.DomainSim(pos, goal, 0, 3, x_bounds=np.array([0, -2.0, 2.0, 0], np.float32))
    s = lcb.DomainSim(pos, goal, 0, 2)
    with pytest.raises(lcb.LcError):
        s.set_state(pos=pos)
