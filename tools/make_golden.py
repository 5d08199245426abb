"""Generate tests/golden/*.npz from O-ref (oracle/_ref/oref = the reference's own translation units, oracle/ref/).
Run in the container that has /root/reference (needed only to BUILD oref):  python tools/make_golden.py
Each fixture holds the reference's table iteration order, initial state, sampled per-frame component state and the
final scores of one reference rollout, so the CPU suite can pin O-canon without the reference being present."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oref  # noqa: E402
from learn_comms_b200 import workloads as wl  # noqa: E402

KEYS = ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")
# (name, scene, seed, weight seed or None, comm, frames to keep (None = to the end), sample stride)
CASES = [
    ("scene4_w2", 4, 0, 2, True, None, 97),       # crossing regiments, learned weights: full rollout
    ("scene5_w7", 5, 0, 7, True, 450, 50),      # up to the first contact the reference BVH resolves differently (Q8)
    ("scene4_nocomm", 4, 0, None, False, None, 97),  # nocomm_norender flavour
    ("scene0_zero", 0, 0, None, True, 700, 50),   # zero config == pure TTC baseline, before the first BVH contact
    ("scene0_w1", 0, 0, 1, True, 1100, 50),
    ("scene1_w5", 1, 0, 5, True, 575, 50),
    ("scene12_w3", 12, 0, 3, True, 1, 1),          # random layout: frame 1 only (initial overlaps hit the BVH path)
    # walled scenes (fixtures also hold the wall rectangles and the statics order); "to the end" where the reference's
    # dynamic BVH never resolves a contact differently from the exact leaf predicate (quirk Q8), else up to that frame
    ("walled_rollout_6_zero", 6, 0, None, True, None, 97),    # curbed intersection, 11 replans
    ("walled_rollout_18_zero", 18, 0, None, True, None, 97),  # two-door, 27 replans
    ("walled_rollout_10_w30", 10, 0, 30, True, None, 97),     # clogged doorway 5 vs 35, 40 replans
    ("walled_rollout_13_w33", 13, 0, 33, True, None, 97),     # hall
    ("walled_rollout_8_w28", 8, 0, 28, True, None, 50),       # doorway 2 vs 2
    ("walled_rollout_3_w23", 3, 0, 23, True, 600, 50),        # pillar: 283 wall contacts before the first Q8 event
    ("walled_rollout_9_w29", 9, 0, 29, True, 1000, 50),       # clogged doorway 20 vs 20
    ("walled_rollout_2_zero", 2, 0, None, True, 2000, 100),   # 60 random posts (walls drawn from the seeded stream)
    ("walled_rollout_11_w31", 11, 0, 31, True, 1600, 100),    # escaping room
]


# comm::x feature-set variants (the #defines inside comm.cpp:22-28): (name, scene, weight seed, O-ref build, LC_FLAG_FEAT_* bits)
VARIANTS = [("scene4_w2_noprox_nosoftsign", 4, 2, "OREF_NOPROX_NOSOFTSIGN", 64 | 128),
            ("scene8_w28_noprox_nosoftsign", 8, 28, "OREF_NOPROX_NOSOFTSIGN", 64 | 128)]


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    only = sys.argv[1:]
    cases = [c + (None, 0) for c in CASES] + [(n, sc, 0, ws, True, None, 97, getattr(_oref, exe), feat) for n, sc, ws, exe, feat in VARIANTS]
    for name, scene, seed, wseed, comm, upto, stride, exe, feat in cases:
        if only and not any(name.startswith(o) for o in only):
            continue
        m_cv, m_f = (None, None) if wseed is None else wl.weights(wseed)
        r = _oref.run(scene, seed, m_cv, m_f, comm=comm, exe=exe)
        frames = [fr for fr in r["frames"] if (upto is None or fr["frame"] <= upto)]
        keep = [fr for fr in frames if fr["frame"] % stride == 0 or fr["frame"] == 1 or fr is frames[-1]]
        d = dict(scene=scene, seed=seed, wseed=-1 if wseed is None else wseed, comm=int(comm), order=r["order"],
                 pos0=r["pos0"], goal=r["goal"], min_time=r["min_time"], frames=np.array([fr["frame"] for fr in keep]),
                 frames_executed=r["frames_executed"], complete=int(upto is None),
                 scores=np.array([r["scores"][k] for k in ("avg_velocity", "group_time", "run_avg", "run_std")], np.float32),
                 n_done=r["scores"]["n_done"], result_file=np.array(r["result_file"]),
                 progress=np.array([[fr["run_avg"], fr["run_std"], fr["avg_velocity"], fr["total_num_done"]] for fr in keep]))
        if r["n_walls"]:
            d["walls"], d["wall_order"] = r["walls"], r["wall_order"]
        if feat:
            d["features"] = feat
        if m_cv is not None:
            d["m_cv"], d["m_f"] = m_cv, m_f
        for k in KEYS:
            d[k] = np.stack([fr[k] for fr in keep])
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **d)
        print(name, "frames kept", len(keep), "of", r["frames_executed"], "scores", r["scores"])


if __name__ == "__main__":
    main()
