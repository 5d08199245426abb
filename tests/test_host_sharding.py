"""Host-side multi-GPU logic on CPU: scenario sharding and the score gather, world_size 2 over gloo (no GPU needed).
The device work itself is replaced by a deterministic stand-in here; the -m gpu tests cover the kernels."""
import os
import socket

import numpy as np
import pytest

from learn_comms_b200 import pso


def test_shard_range_partitions():
    for n in (0, 1, 7, 64, 4096, 4099):
        for world in (1, 2, 3, 8):
            blocks = [pso.shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[r][1] == blocks[r + 1][0] for r in range(world - 1))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1


def test_config_split_and_files(tmp_path):
    cfg = np.random.RandomState(0).uniform(-1, 1, pso.M_SHAPE)
    m_cv, m_f = pso.split_config(cfg)
    assert m_cv.shape == (3, 8) and m_f.shape == (2, 3)
    assert np.array_equal(m_cv, cfg[:3].astype(np.float32)) and np.array_equal(m_f, cfg[3:, :3].astype(np.float32))
    p = tmp_path / "comms.config"
    with open(p, "w") as f:
        pso.write_config(f, cfg)
    lines = p.read_text().splitlines()
    assert len(lines) == 5 and all(len(l.split(" ")) == 8 for l in lines)
    # byte-compatible with the product's parser (== comm::init, comm.cpp:80-121)
    import learn_comms_b200 as lcb
    a, b = lcb.read_config(str(p))
    assert np.array_equal(a, m_cv) and np.array_equal(b, m_f)
    r = tmp_path / "comms.result"
    r.write_text("0.5\nnan\n")
    assert pso.read_result(str(r)) == 1e99
    r.write_text("0.5\n12.25\n")
    assert pso.read_result(str(r)) == 12.25


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_items, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = pso.shard_range(n_items, rank, world)
    idx = np.arange(lo, hi)
    local = dict(avg_velocity=(idx * 0.5).astype(np.float32), group_time=(idx * 3 + 1).astype(np.float32),
                 n_done=(idx % 64).astype(np.int32), frames=(idx + 600).astype(np.int32))
    out = pso.gather_scores(local, n_items, rank, world, dist)
    q.put((rank, {k: v.tolist() for k, v in out.items()}, {k: str(v.dtype) for k, v in out.items()}))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_items", [9, 64])
def test_score_gather_world2_gloo(n_items):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_items, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    idx = np.arange(n_items)
    for rank, out, dtypes in got:
        assert out["avg_velocity"] == (idx * 0.5).astype(np.float32).tolist()
        assert out["group_time"] == (idx * 3 + 1).astype(np.float32).tolist()
        assert out["n_done"] == (idx % 64).tolist() and out["frames"] == (idx + 600).tolist()
        assert dtypes["n_done"] == "int32" and dtypes["avg_velocity"] == "float32"


def test_domain_score_merge_is_pure_host_logic():
    """physics.cpp:72-76 across ranks: the whole crowd's avg_velocity is the speed of the highest-id not-done agent over
    all ranks (lastvel key = ((id + 1) << 32) | speed bits); non-finite speeds elsewhere poison it (NaN * 0)."""
    import learn_comms_b200 as lcb
    bits = lambda f: int(np.array([f], np.float32).view(np.uint32)[0])
    r0 = dict(avg_velocity=0.25, group_time=3.0, run_avg=1.0, run_std=0.5, n_done=7)
    key = lambda i, sp: ((i + 1) << 32) | bits(sp)
    # nobody left that is not done: the previous frame's value stands
    assert lcb.DomainSim.merge_scores(r0, [(0, 0, 0.25), (0, 0, 0.25)])["avg_velocity"] == 0.25
    # highest id wins, whichever rank holds it
    out = lcb.DomainSim.merge_scores(r0, [(key(10, 0.5), 0, 0.25), (key(99, 0.75), 0, 0.25), (key(50, 0.1), 0, 0.25)])
    assert out["avg_velocity"] == 0.75 and out["group_time"] == 3.0 and out["n_done"] == 7
    # a non-finite speed that is not the last one poisons the chain; the last one being infinite is kept as is
    assert np.isnan(lcb.DomainSim.merge_scores(r0, [(key(10, 0.5), 1, 0.25), (key(99, 0.75), 0, 0.25)])["avg_velocity"])
    assert np.isinf(lcb.DomainSim.merge_scores(r0, [(key(10, 0.5), 0, 0.25), (key(99, np.inf), 1, 0.25)])["avg_velocity"])
    assert np.isnan(lcb.DomainSim.merge_scores(r0, [(key(10, 0.5), 0, np.nan), (key(99, 0.75), 0, np.nan)])["avg_velocity"])


def test_dd_strips_equalise_the_local_counts():
    """lc_dd_strips (host code; lc_dd_create's default cuts): the LOCAL counts — own strip plus ghost bands — are what the
    ranks' kernels work on, so those are equalised, not the owned counts: interior strips own fewer agents than edge strips."""
    import learn_comms_b200 as lcb
    rs = np.random.RandomState(11)
    n, band = 200_000, 11.9
    pos = np.stack([rs.uniform(-1000, 1000, n), rs.uniform(-1000, 1000, n)], 1).astype(np.float32)
    xs = np.sort(pos[:, 0])
    for world in (2, 4, 8):
        xb = lcb.dd_strips(pos, world, band)
        assert xb[0] < -1e38 and xb[world] > 1e38 and (np.diff(xb[1:world]) > 0).all()
        lo = np.where(np.arange(world) == 0, -np.inf, xb[:world] - band)
        hi = np.where(np.arange(world) == world - 1, np.inf, xb[1:] + band)
        local = np.array([np.searchsorted(xs, h) - np.searchsorted(xs, l) for l, h in zip(lo, hi)])
        owned = np.array([np.searchsorted(xs, xb[r + 1]) - np.searchsorted(xs, xb[r]) for r in range(world)])
        assert owned.sum() == n
        assert local.max() - local.min() <= 0.002 * local.mean(), (world, local)
        if world > 2:   # the quantile cuts would leave the interior strips a ghost band heavier
            assert owned[0] > owned[1] and owned[-1] > owned[-2]
            q = np.array([np.searchsorted(xs, np.inf if r == world - 1 else xs[(r + 1) * n // world] + band) -
                          np.searchsorted(xs, -np.inf if r == 0 else xs[r * n // world] - band) for r in range(world)])
            assert q.max() - q.min() > 10 * (local.max() - local.min())
    # a lopsided crowd: three quarters of the agents in the left tenth
    pos2 = pos.copy()
    pos2[: 3 * n // 4, 0] = rs.uniform(-1000, -800, 3 * n // 4)
    xb = lcb.dd_strips(pos2, 4, band)
    xs2 = np.sort(pos2[:, 0])
    local = [np.searchsorted(xs2, (np.inf if r == 3 else xb[r + 1] + band)) - np.searchsorted(xs2, (-np.inf if r == 0 else xb[r] - band)) for r in range(4)]
    assert max(local) - min(local) <= 0.01 * np.mean(local), local
    with pytest.raises(lcb.LcError):
        lcb.dd_strips(pos[:5], 4, band)   # fewer than two agents per strip
