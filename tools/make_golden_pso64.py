"""PSO-fitness agreement fixture: 64 random particles (5 x 8 comms.config matrices drawn as
pso_opt.py:76-83 does, U(-1, 1) on the 30 live entries) evaluated by the REFERENCE'S OWN translation units (O-ref) on
scenes 0, 4, 5, 12 (seed 0), plus the all-zero baseline of pso_opt.py:312-319.  Stores the configs, each comms.result
fitness (pso_opt.py:26-41) and the frames executed in tests/golden/pso64_oref.npz; tests/test_ensemble.py computes the
loss of pso_opt.py:156-194 from the product's scores and compares ranking and argmin.
Run where /root/reference is present (O-ref must be built):  python tools/make_golden_pso64.py"""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oref  # noqa: E402

SCENES = (0, 4, 5, 12)
P = 64


def particles(p=P, seed=64):
    rs = np.random.RandomState(seed)
    cfg = rs.uniform(-1.0, 1.0, (p, 5, 8))
    cfg[:, 3:, 3:] = 0.0  # pso_opt.py:78-80: entries redundant to TTC stay zero
    return cfg


def one(job):
    k, scene, cfg = job
    c32 = cfg.astype(np.float32)
    r = _oref.run(scene, 0, c32[:3], c32[3:, :3], dump_every=0)
    gt = float(r["result_file"][1])  # what read_result parses from comms.result line 2
    if np.isnan(gt) or gt > 1e99:
        gt = 1e99
    return [gt, r["frames_executed"], float(r["result_file"][0])]


def main():
    global SCENES, P
    name = "pso64_oref.npz"
    if len(sys.argv) > 1 and sys.argv[1] == "walled":  # the walled half of pso_opt.py's default scene universe (pso_opt.py:514)
        SCENES, P, name = (8, 9, 13, 18), 32, "pso32_walled_oref.npz"
    cfg = particles(P, 64 if name.startswith("pso64") else 32)
    jobs = [(k, s, cfg[k]) for k in range(P) for s in SCENES] + [(-1, s, np.zeros((5, 8))) for s in SCENES]
    with ThreadPoolExecutor(max_workers=os.cpu_count()) as ex:
        out = np.array(list(ex.map(one, jobs)), np.float64)
    res, base = out[:P * len(SCENES)].reshape(P, len(SCENES), 3), out[P * len(SCENES):]
    path = os.path.join(ROOT, "tests", "golden", name)
    np.savez_compressed(path, scenes=np.array(SCENES, np.int32), configs=cfg, fitness=res[:, :, 0], frames=res[:, :, 1].astype(np.int32),
                        avg_velocity=res[:, :, 2], baseline=base[:, 0], baseline_frames=base[:, 1].astype(np.int32))
    loss = (res[:, :, 0] / base[None, :, 0]).sum(1)
    print("baseline", base[:, 0], "loss min", loss.min(), "argmin", int(loss.argmin()), "median", np.median(loss), "max", loss.max())


if __name__ == "__main__":
    main()
