"""Dev helper: run scene rollouts on the GPU and in O-canon side by side and print where they diverge."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import learn_comms_b200 as lcb
import _oracle


def cmp_state(a, b, tag):
    ok = True
    for k in ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done"):
        x, y = a[k], b[k]
        same = np.array_equal(x, y, equal_nan=True) if x.dtype.kind == "f" else np.array_equal(x, y)
        if not same:
            d = np.abs(x.astype(np.float64) - y.astype(np.float64))
            print(f"  {tag}: {k} differs: max abs {np.nanmax(d):.3e} at {np.unravel_index(np.nanargmax(d), d.shape)} ({(d>0).sum()} elems)")
            ok = False
    return ok


def run(scene_id, weights_seed=None, frames=None, grid=False):
    pos, goal, md, rng = lcb.scene(scene_id, 0)
    m_cv = m_f = None
    if weights_seed is not None:
        r = np.random.RandomState(weights_seed)
        m_cv = r.uniform(-1, 1, (3, 8)).astype(np.float32)
        m_f = r.uniform(-1, 1, (2, 3)).astype(np.float32)
    g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, debug_lists=True)
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, grid=grid)
    o.record_neighbors(True)
    f = 0
    bad = 0
    while not o.finished and (frames is None or f < frames):
        g.step(1); o.step(1); f += 1
        ok = cmp_state(g.state(), o.state(), f"scene {scene_id} frame {f}")
        if not np.array_equal(g.partners(), o.partners()):
            print(f"  frame {f}: partners differ", np.nonzero(g.partners() != o.partners())[0][:10]); ok = False
        if not np.array_equal(g.neighbor_counts(), o.neighbor_counts()):
            print(f"  frame {f}: neighbour counts differ"); ok = False
        else:
            gp, gi = g.neighbor_lists(); op, oi = o.neighbor_lists()
            if not (np.array_equal(gp, op) and np.array_equal(gi, oi)):
                print(f"  frame {f}: neighbour lists differ"); ok = False
        cg = g.counters()
        if cg["overlap_pairs"] != o.overlap_pairs or cg["collision_hits"] != o.collision_hits:
            print(f"  frame {f}: counters differ gpu={cg} oracle=({o.overlap_pairs},{o.collision_hits})"); ok = False
        if not ok:
            bad += 1
            if bad >= 3:
                break
    print(f"scene {scene_id} w={weights_seed}: {f} frames, gpu finished={g.finished} frame={g.frame}")
    print("  gpu   ", g.scores())
    print("  oracle", o.scores())
    return bad == 0


if __name__ == "__main__":
    allok = True
    for sc, ws in ((0, None), (0, 1), (4, 2), (12, 3)):
        allok &= run(sc, ws)
    print("ALL OK" if allok else "MISMATCH")
