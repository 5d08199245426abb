"""In-process batched replacement of pso_opt.py's fitness evaluation (pso_opt.py:236-267 simulate_set / simulate).

The reference evaluates one particle on one scene by writing `comms.config`, fork-exec'ing `comm_norender <scene> --seed S
--data D`, and reading `comms.result` back (about 40 x S process launches per PSO iteration).  Here every
(particle, scene) pair of an iteration is one scenario of ONE lc_batch_run launch per GPU; the file formats are kept
(write_config / read_result below are byte-compatible with pso_opt.py:17-41) so archives stay readable by the reference's
scripts.  Multi-GPU: scenarios are independent, so each rank evaluates a contiguous block and only the scores
(avg_velocity, group_time, n_done, frames: 16 B per scenario) are gathered.
"""
import numpy as np

from . import api

N_COMM = 3
M_SHAPE = (N_COMM + 2, N_COMM + 5)  # pso_opt.py:13-14: the 5 x 8 comms.config matrix


def write_config(file, config):
    """pso_opt.py:17-23: 5 lines of 8 space-separated str(float) tokens."""
    for i in range(M_SHAPE[0]):
        file.write(" ".join(str(config[i][j]) for j in range(M_SHAPE[1])))
        file.write("\n")


def aggregate(avg_vel, confident_time):
    """pso_opt.py:30-35: the fitness is the mean + 3 sigma overhead time; NaN or absurd values become 1e99."""
    if np.isnan(confident_time) or float(confident_time) > 1e99:
        return 1e99
    return confident_time


def read_result(result_file):
    """pso_opt.py:26-41."""
    with open(result_file, "r") as f:
        avg_vel = float(f.readline().strip())
        confident_time = float(f.readline().strip())
    return aggregate(avg_vel, confident_time)


def split_config(config):
    """5 x 8 comms.config matrix -> (M_cv 3 x 8, M_f 2 x 3) exactly as comm::init reads it (comm.cpp:80-131):
    rows 0-2 are M_c | M_vy M_s M_py M_d M_g, the first three entries of rows 3-4 are M_f."""
    config = np.asarray(config, np.float32)
    assert config.shape == M_SHAPE
    return np.ascontiguousarray(config[:3]), np.ascontiguousarray(config[3:, :3])


def shard_range(n_items, rank, world):
    """Contiguous block of items for `rank` (sizes differ by at most one; empty blocks only if world > n_items)."""
    base, rem = divmod(int(n_items), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_scores(local, n_items, rank, world, dist):
    """Assemble per-scenario score arrays from the per-rank blocks.  `dist` is torch.distributed (gloo or nccl) or
    None.  Every rank returns the full arrays."""
    if dist is None or world == 1:
        return local
    import torch
    keys = sorted(local)
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    sizes = [shard_range(n_items, r, world) for r in range(world)]
    width = max(hi - lo for lo, hi in sizes)
    mine = torch.zeros((len(keys), width), dtype=torch.float64, device=dev)
    lo, hi = sizes[rank]
    for k, key in enumerate(keys):
        mine[k, :hi - lo] = torch.from_numpy(np.asarray(local[key], np.float64)).to(dev)
    parts = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    out = {}
    for k, key in enumerate(keys):
        full = np.concatenate([parts[r][k, :sizes[r][1] - sizes[r][0]].cpu().numpy() for r in range(world)])
        out[key] = full.astype(np.asarray(local[key]).dtype)
    return out


class BatchEvaluator:
    """Fixed set of scenario layouts (pos, goal, max_duration, rng per scenario), re-evaluated with new per-scenario
    parameter matrices every PSO iteration.  evaluate() returns full-length host arrays on every rank."""

    def __init__(self, pos, goal, max_duration, rng_state, device=0, rank=0, world=1, dist=None, comm=True, order="reference",
                 walls=None):
        """order: table iteration order of every scenario (lc_batch_desc::table_order).  Default = the reference's own
        (quirk Q4), which is what the comm_norender processes pso_opt.py spawns would use.  walls [W][4]: the wall layout
        all scenarios share (a walled catalogue scene evaluated for many particles)."""
        self.S, self.n = int(pos.shape[0]), int(pos.shape[1])
        self.rank, self.world, self.dist = rank, world, dist
        self.lo, self.hi = shard_range(self.S, rank, world)
        sl = slice(self.lo, self.hi)
        self.batch = None
        if self.hi > self.lo:
            S_loc = self.hi - self.lo
            self.batch = api.Batch(pos[sl], goal[sl], np.zeros((S_loc, 3, 8), np.float32), np.zeros((S_loc, 2, 3), np.float32),
                                   np.asarray(max_duration, np.float32)[sl], np.asarray(rng_state, np.uint32)[sl],
                                   comm=comm, device=device, order=order, walls=walls)
        S_loc = self.hi - self.lo
        self.h2d_bytes = S_loc * (30 * 4 + self.n * (16 + 8 + 8 + 16 + 16))  # weights + the initial tables reset re-uploads
        self.d2h_bytes = S_loc * 56                                          # one control block per scenario

    def evaluate(self, m_cv, m_f):
        """m_cv [S][3][8], m_f [S][2][3] (host) -> dict of [S] arrays: avg_velocity, group_time, n_done, frames, fitness."""
        sl = slice(self.lo, self.hi)
        if self.batch is not None:
            self.batch.reset(np.asarray(m_cv, np.float32)[sl], np.asarray(m_f, np.float32)[sl])
            local = self.batch.run()
        else:
            local = dict(avg_velocity=np.zeros(0, np.float32), group_time=np.zeros(0, np.float32),
                         n_done=np.zeros(0, np.int32), frames=np.zeros(0, np.int32))
        out = gather_scores(local, self.S, self.rank, self.world, self.dist)
        out["fitness"] = np.array([aggregate(a, g) for a, g in zip(out["avg_velocity"], out["group_time"])], np.float64)
        return out


def simulate_set(scene_layouts, configs, device=0, rank=0, world=1, dist=None, comm=True, order="reference"):
    """pso_opt.py:236-242 simulate_set for MANY particles at once: scene_layouts = list of (pos[n,2], goal[n,2],
    max_duration, rng_state[, walls[W,4]]) — lc_scene_full output for the reference's scenes, walled ones included —
    configs = [P] 5 x 8 matrices.  One batch per scene (a batch holds scenarios of one agent count and one wall layout),
    every particle a scenario of it.  Returns fitness [P][len(scene_layouts)] (read_result's aggregate of each
    comms.result) and the raw per-scene score arrays."""
    L, P = len(scene_layouts), len(configs)
    split = [split_config(c) for c in configs]
    m_cv = np.stack([sp[0] for sp in split])
    m_f = np.stack([sp[1] for sp in split])
    fitness = np.zeros((P, L), np.float64)
    raw = []
    for l, lay in enumerate(scene_layouts):
        pos0, goal0, md0, rng0 = lay[:4]
        walls = lay[4] if len(lay) > 4 else None
        pos = np.broadcast_to(np.asarray(pos0, np.float32), (P,) + pos0.shape).copy()
        goal = np.broadcast_to(np.asarray(goal0, np.float32), (P,) + goal0.shape).copy()
        ev = BatchEvaluator(pos, goal, np.full(P, md0, np.float32), np.full(P, rng0, np.uint32), device=device, rank=rank, world=world,
                            dist=dist, comm=comm, order=order, walls=walls)
        out = ev.evaluate(m_cv, m_f)
        fitness[:, l] = out["fitness"]
        raw.append(out)
    return fitness, raw
