"""The walled half of the rollout-level calibration (see tools/make_golden_ensemble.py): the reference's own spread of
comms.result under a 1-ulp perturbation of the start positions on walled catalogue scenes — planner, sight-line replans,
wall forces and wall push-out all in play.  Writes tests/golden/ensemble_walled_oref.npz.
Run where /root/reference is present (O-ref must be built):  python tools/make_golden_ensemble_walled.py"""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle  # noqa: E402
import _oref  # noqa: E402
from learn_comms_b200 import workloads as wl  # noqa: E402

MEMBERS = 24
CASES = [(3, 23), (8, 28), (9, None), (9, 29), (13, 33), (18, None)]  # (scene, weight seed)


def one(job):
    scene, wseed, k = job
    pos = _oracle.scene_full(scene, 0)[0]
    m_cv, m_f = (None, None) if wseed is None else wl.weights(wseed)
    r = _oref.run(scene, 0, m_cv, m_f, dump_every=0, pos_override=_oref.perturb_ulp(pos, k))
    sc = r["scores"]
    return [sc["avg_velocity"], sc["group_time"], r["frames_executed"], sc["n_done"]]


def main():
    jobs = [(s, w, k) for (s, w) in CASES for k in range(MEMBERS)]
    with ThreadPoolExecutor(max_workers=os.cpu_count()) as ex:
        out = list(ex.map(one, jobs))
    res = np.array(out, np.float64).reshape(len(CASES), MEMBERS, 4)
    path = os.path.join(ROOT, "tests", "golden", "ensemble_walled_oref.npz")
    np.savez_compressed(path, cases=np.array([(s, -1 if w is None else w) for s, w in CASES], np.int32), members=MEMBERS, result=res)
    for (s, w), r in zip(CASES, res):
        gt = r[:, 1]
        print(f"scene {s:2d} weights {w}: group_time mean {gt.mean():.4f} sd {gt.std(ddof=1):.4f} min {gt.min():.4f} max {gt.max():.4f} frames {r[:, 2].mean():.0f}")


if __name__ == "__main__":
    main()
