"""Generate tests/golden/walled_scene3.npz from O-ref (oracle/_ref/oref): the reference's own rollout of scene 3 (60
agents on a circle, one 4 m pillar, demo.cpp:176-187) with learned weights, every frame of the first FRAMES: positions,
goal_dist, num_done, local_goal, plus table order, initial state and min_time.  tests/test_planner.py replays the
waypoint bookkeeping over it.  Run where /root/reference is present:  python tools/make_golden_walled.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oref  # noqa: E402
from learn_comms_b200 import workloads as wl  # noqa: E402

FRAMES = 700  # covers the first lost_frames replans (frame 361 on) and an unreachable replan (frame 666)


def main():
    m_cv, m_f = wl.weights(4)
    r = _oref.run(3, 0, m_cv, m_f, dump_every=1, max_frames=FRAMES)
    fr = r["frames"][:FRAMES]
    d = dict(n=r["n"], order=r["order"], pos0=r["pos0"], goal=r["goal"], min_time=r["min_time"],
             frame=np.array([f["frame"] for f in fr], np.int32), pos=np.stack([f["pos"] for f in fr]),
             goal_dist=np.stack([f["goal_dist"] for f in fr]), num_done=np.stack([f["num_done"] for f in fr]).astype(np.int8),
             local_goal=np.stack([f["local_goal"] for f in fr]))
    path = os.path.join(ROOT, "tests", "golden", "walled_scene3.npz")
    np.savez_compressed(path, **d)
    print("frames", len(fr), "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
