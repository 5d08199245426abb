"""Teacher-forced replay of the waypoint bookkeeping of a walled reference rollout.  TEST INFRASTRUCTURE ONLY.
Pieces under test: the product's host planner (lc_planner_*, stateful roadmap, reference plan order), the sight-line
gate and replan rule of ai::update_agents (ai.cpp:146-166: lost_frames >= 360, at most two replans per frame in table
order, replans start from the agent's current position), the no-plan branch of calc_sum_force (LMP.cpp:259-270) and
LMP::lookahead (O-canon's canon_lookahead)."""
import numpy as np

import _oracle
import _oref
import learn_comms_b200 as lcb

SCENE3_WALLS = np.array([[0, 0, 4, 4]], np.float32)  # demo.cpp:176-187: one 4 m pillar at the centre


def follow(r, walls, radius=0.15):
    """Teacher-forced replay of the plan bookkeeping of ai::update_agents (ai.cpp:146-166) + LMP::lookahead over a
    reference rollout: returns the first frame at which num_done / local_goal differ, or None."""
    L = _oracle.lib()
    walls = np.ascontiguousarray(walls, np.float32)
    P = lcb.Planner(walls, radius)
    n, order, fr = r["n"], r["order"], r["frames"]
    plans = [None] * n
    for i in order:
        plans[i], _ = P.plan(r["pos0"][i], r["goal"][i])
    lost = np.zeros(n, np.int64)
    replans = 0
    prev = dict(pos=r["pos0"], num_done=np.zeros(n, np.int32), local_goal=np.zeros((n, 2), np.float32))
    for f in fr:
        replanned = 0
        rec = np.zeros(n, _oref.LOOK_REC)
        empty = np.zeros(n, bool)
        nd_keep = np.zeros(n, np.int32)
        for i in order:
            pos = np.ascontiguousarray(prev["pos"][i], np.float32)
            nd, plan = int(prev["num_done"][i]), plans[i]
            has_plan = len(plan) > 0
            visible = False
            if has_plan and nd < len(plan):
                tgt = np.ascontiguousarray(plan[nd], np.float32)
                visible = bool(L.canon_cspace_line_of_sight(walls.ctypes.data, len(walls), radius, pos.ctypes.data, tgt.ctypes.data))
            lost[i] = 0 if visible else lost[i] + 1
            done = has_plan and nd == len(plan)
            need = (not has_plan) or (done and f["goal_dist"][i] > 1.0) or ((not done) and (not visible) and lost[i] >= 360)
            if replanned < 2 and need:
                plans[i], _ = P.plan(pos, r["goal"][i])
                plan, nd = plans[i], 0
                replanned += 1
                replans += 1
                lost[i] = 0
            if len(plan) == 0:  # LMP.cpp:259-270: no plan -> local_goal = own position, no lookahead
                empty[i] = True
                nd_keep[i] = nd
                rec[i] = (pos, pos, 0, 1, np.zeros((8, 2), np.float32))
                continue
            assert len(plan) <= 8
            rec[i] = (pos, prev["local_goal"][i], nd, len(plan), np.vstack([plan, np.zeros((8 - len(plan), 2), np.float32)]))
        out = _oracle.lookahead(walls, radius, rec)
        out["num_done"][empty] = nd_keep[empty]
        out["lg"][empty] = rec["pos"][empty]
        if not (np.array_equal(out["num_done"], f["num_done"]) and np.array_equal(out["lg"], f["local_goal"])):
            return f["frame"], replans
        prev = f
    return None, replans
