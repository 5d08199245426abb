"""Generate tests/golden/planner.npz from oracle/_ref/oref_plan (the reference's own Cspace2d / PRM / GMP::plan_one, see
oracle/ref/plan_main.cpp).  Run in the container that has /root/reference (needed only to BUILD the harness):
make -C oracle/ref && python tools/make_golden_planner.py
The fixture holds seeded wall layouts, (start, goal) pairs and the reference's waypoint lists, roadmap sizes and
min_time values, so the CPU suite can pin the product's host-side planner (lc_planner_*) without the reference."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oref  # noqa: E402

MAX_WP = 32


def layout(seed, n_agents=40):
    """Three families: a row of 0.3 m tiles with a doorway (wall_scale 0.3, demo.cpp:164), scattered 4 m blocks
    (wall_scale 4.0, demo.cpp:180), arbitrary rectangles.  Starts and goals are uniform in the 22 m square, so some
    lie inside an inflated wall (unreachable -> empty plan, which has to match too)."""
    g = np.random.default_rng(seed)
    kind = seed % 3
    if kind == 0:
        xs = np.arange(-6, 6.01, 0.3)
        xs = xs[np.abs(xs - g.uniform(-3, 3)) > 0.9]
        walls = [[x, 0, 0.3, 0.3] for x in xs]
    elif kind == 1:
        walls = [[x, y, 4, 4] for x in (-6, 0, 6) for y in (-6, 0, 6) if g.random() < 0.6]
    else:
        walls = [[g.uniform(-9, 9), g.uniform(-9, 9), g.uniform(0.3, 5), g.uniform(0.3, 5)] for _ in range(g.integers(1, 9))]
    start = g.uniform(-11, 11, (n_agents, 2)).astype(np.float32)
    goal = g.uniform(-11, 11, (n_agents, 2)).astype(np.float32)
    return np.array(walls, np.float32).reshape(-1, 4), start, goal


def look_records(seed, start, plans, per_plan=6):
    """Agent positions scattered around their own path (so waypoints get reached, skipped or hidden behind walls), with
    every admissible num_done, for LMP::lookahead.  Plans longer than 8 waypoints or empty are left out."""
    g = np.random.default_rng(1000 + seed)
    recs = []
    for i, (_, wp) in enumerate(plans):
        if not 1 <= len(wp) <= 8:
            continue
        line = np.vstack([start[i][None], wp])
        for _ in range(per_plan):
            k = g.integers(0, len(line) - 1)
            t = g.random()
            pos = line[k] * (1 - t) + line[k + 1] * t + g.normal(0, 0.4, 2)
            r = np.zeros((), _oref.LOOK_REC)
            r["pos"], r["lg"] = pos, g.uniform(-5, 5, 2)
            r["num_done"], r["n_plan"] = g.integers(0, len(wp) + 1), len(wp)
            r["plan"][:len(wp)] = wp
            recs.append(r)
    return np.array(recs, _oref.LOOK_REC)


def main():
    d = {}
    seeds = list(range(12))
    for s in seeds:
        walls, start, goal = layout(s)
        nv, ne, plans = _oref.run_plan(walls, 0.15, start, goal)
        wp = np.zeros((len(start), MAX_WP, 2), np.float32)
        cnt = np.zeros(len(start), np.int32)
        for i, (_, p) in enumerate(plans):
            assert len(p) <= MAX_WP
            wp[i, :len(p)] = p
            cnt[i] = len(p)
        d.update({f"walls{s}": walls, f"start{s}": start, f"goal{s}": goal, f"wp{s}": wp, f"cnt{s}": cnt,
                  f"min_time{s}": np.array([p[0] for p in plans], np.float32), f"roadmap{s}": np.array([nv, ne], np.int32)})
        recs = look_records(s, start, plans)
        d[f"look_in{s}"], d[f"look_out{s}"] = recs, _oref.run_look(walls, 0.15, recs)
        print("layout", s, "walls", len(walls), "roadmap", nv, ne, "waypoints", np.bincount(cnt))
    d["seeds"] = np.array(seeds, np.int32)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "planner.npz"), **d)


if __name__ == "__main__":
    main()
