"""Synthetic inputs of the BASELINE.json configs.  Host-side numpy only; float32 throughout.

Every generator returns (pos[n,2], goal[n,2], max_duration, rng_state) in the conventions of lc_desc / demo::init
(demo.cpp:611-648).  Placement follows the reference's own scene helpers where one exists:
circle = make_radial_bots + opposite_goal (demo.cpp:48-54,30-32), corridor = make_regimented_bots + mirror_x_goal
(demo.cpp:72-95,33-35), plaza = scene-12-style uniform draws (demo.cpp:356-373) on a jittered lattice so that no
two agents start overlapped.
"""
import numpy as np

F = np.float32


def lcg_seed(seed):
    """Seeder::seed (Seeder.cpp:15-18) -> minstd_rand0 state."""
    x = int(seed) % 2147483647
    return 1 if x == 0 else x


def circle(n=1000, radius=80.0, seed=1):
    """C2: n agents on a circle, antipodal goals (demo.cpp:48-54 generalised; 1k @ r=80 -> 0.50 m spacing)."""
    i = np.arange(n, dtype=F)
    radian = i / F(n) * F(2.0) * F(np.pi)
    pos = np.stack([F(radius) * np.cos(radian, dtype=F), F(radius) * np.sin(radian, dtype=F)], 1).astype(F)
    goal = (-pos).astype(F)
    # longest path is the diameter at unit speed; leave 50 % slack like scene 0 (20 m -> 45 s)
    return pos, goal, float(2.25 * 2 * radius), lcg_seed(seed)


def corridor(cols=500, rows=100, gap=20.0, spacing=1.0, seed=4):
    """C4: two cols x rows blocks on a lattice facing each other along x, mirror_x goals (demo.cpp:33-35,72-95).
    Defaults give 100 000 agents in a ~1 020 m x 100 m band; each block walks towards the other's start."""
    xs = (np.arange(cols, dtype=F) * F(spacing) + F(gap / 2)).astype(F)
    ys = ((np.arange(rows, dtype=F) - F((rows - 1) / 2)) * F(spacing)).astype(F)
    gx, gy = np.meshgrid(xs, ys, indexing="xy")
    right = np.stack([gx.ravel(), gy.ravel()], 1)
    left = right * np.array([-1, 1], F)
    pos = np.concatenate([left, right]).astype(F)
    goal = (pos * np.array([-1, 1], F)).astype(F)
    return pos, goal, float(2.25 * 2 * (cols * spacing + gap)), lcg_seed(seed)


def plaza(n=1_000_000, side=2000.0, seed=5, goal_range=None):
    """C5: n agents in a side x side square: jittered lattice (pitch = side/sqrt(n), jitter keeps >= 0.35 m gaps),
    goals uniform in the plaza (scene-12 style, demo.cpp:362-371) or within goal_range metres of the start."""
    rs = np.random.RandomState(seed)
    m = int(np.ceil(np.sqrt(n)))
    pitch = side / m
    jit = max(0.0, (pitch - 0.35) / 2)
    idx = rs.permutation(m * m)[:n]
    cx = (idx % m).astype(np.float64) * pitch + pitch / 2 - side / 2
    cy = (idx // m).astype(np.float64) * pitch + pitch / 2 - side / 2
    pos = np.stack([cx + rs.uniform(-jit, jit, n), cy + rs.uniform(-jit, jit, n)], 1).astype(F)
    if goal_range is None:
        goal = rs.uniform(-side / 2, side / 2, (n, 2)).astype(F)
    else:
        goal = (pos + rs.uniform(-goal_range, goal_range, (n, 2))).astype(F)
    return pos, goal, 1.0e9, lcg_seed(seed)


def weights(seed, scale=1.0, batch=None):
    """30 live comms.config parameters ~ U(-scale, scale) (pso_opt.py:76-83,325-326): M_cv 3x8 and M_f 2x3."""
    rs = np.random.RandomState(seed)
    if batch is None:
        return rs.uniform(-scale, scale, (3, 8)).astype(F), rs.uniform(-scale, scale, (2, 3)).astype(F)
    return rs.uniform(-scale, scale, (batch, 3, 8)).astype(F), rs.uniform(-scale, scale, (batch, 2, 3)).astype(F)


def pso_layouts(n_agents=64):
    """C3: the 8 wall-free 64-agent layouts: circles R=10/13, 4-way crossing with axial and with
    antipodal goals (demo.cpp:96-104,189-203 with 16 per regiment), 4 random layouts in [-10,10]^2 (demo.cpp:358-373)."""
    out = []
    for r in (10.0, 13.0):
        p, g, _, _ = circle(n_agents, r)
        out.append((p, g))
    per = n_agents // 4
    colsn = 8 if per % 8 == 0 else 4
    d = 6.0
    blocks = []
    for facing in ((1, 0), (-1, 0), (0, 1), (0, -1)):
        f = np.array(facing, F)
        rightf = np.array([f[1], -f[0]], F)
        head = -f * F(d) - rightf * F((colsn - 1) / 2)
        k = np.arange(per)
        p = head[None] + rightf[None] * (k % colsn)[:, None].astype(F) + (-f)[None] * (k // colsn)[:, None].astype(F)
        blocks.append(p.astype(F))
    p = np.concatenate(blocks).astype(F)
    ax = np.where(np.abs(p[:, 0]) - np.abs(p[:, 1]) >= 0, 1.0, -1.0).astype(F)
    out.append((p, np.stack([-ax * p[:, 0], ax * p[:, 1]], 1).astype(F)))  # axial_goal, demo.cpp:36-41
    out.append((p, (-p).astype(F)))
    for s in range(4):
        rs = np.random.RandomState(100 + s)
        # rejection to >= 0.4 m so no layout starts overlapped
        pts = []
        while len(pts) < n_agents:
            c = rs.uniform(-10, 10, 2)
            if all((c[0] - q[0]) ** 2 + (c[1] - q[1]) ** 2 >= 0.16 for q in pts):
                pts.append(c)
        out.append((np.array(pts, F), rs.uniform(-10, 10, (n_agents, 2)).astype(F)))
    return out


def pso_batch(n_scenarios=4096, n_agents=64, seed=3, scale=1.0):
    """C3: n_scenarios = (n_scenarios/8 weight vectors) x 8 layouts; per-scenario comm parameters, 45 s limit."""
    lay = pso_layouts(n_agents)
    L = len(lay)
    P = (n_scenarios + L - 1) // L
    m_cv_p, m_f_p = weights(seed, scale, batch=P)
    pos = np.zeros((n_scenarios, n_agents, 2), F)
    goal = np.zeros((n_scenarios, n_agents, 2), F)
    m_cv = np.zeros((n_scenarios, 3, 8), F)
    m_f = np.zeros((n_scenarios, 2, 3), F)
    for s in range(n_scenarios):
        p, l = divmod(s, L)
        pos[s], goal[s] = lay[l]
        m_cv[s], m_f[s] = m_cv_p[p], m_f_p[p]
    md = np.full(n_scenarios, 45.0, F)
    rng = np.full(n_scenarios, 1, np.uint32)
    return pos, goal, m_cv, m_f, md, rng
