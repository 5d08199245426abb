"""tests/golden/catalogue_oref.npz: what the reference's own demo::init (O-ref, oracle/_ref/oref) builds for every scene id of
demo.cpp:141-511 — start positions, goals, wall rectangles in creation order, the iteration order of its agent and wall
tables.  Run in the build container (needs oracle/_ref/oref, i.e. /root/reference at build time):
    python tools/make_golden_catalogue.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oref  # noqa: E402

out = {}
cases = [(s, 0) for s in range(19)] + [(12, 3)]
for scene, seed in cases:
    ref = _oref.run(scene, seed=seed, max_frames=1, dump_every=1)
    k = f"s{scene}_{seed}"
    out[k + "_pos"], out[k + "_goal"], out[k + "_walls"] = ref["pos0"], ref["goal"], ref["walls"]
    out[k + "_order"], out[k + "_wall_order"] = ref["order"].astype(np.int32), ref["wall_order"].astype(np.int32)
out["cases"] = np.array(cases, np.int32)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "catalogue_oref.npz"), **out)
print("wrote", len(cases), "scenes")
