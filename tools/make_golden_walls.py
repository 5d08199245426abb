"""Generate tests/golden/wall_geometry.npz from oracle/_ref/oref_geom (the reference's own BoundVolume.cpp / LMP.cpp /
PRM.cpp, see oracle/ref/geom_main.cpp).  Run in the container that has /root/reference (needed only to BUILD the
harness):  make -C oracle/ref && python tools/make_golden_walls.py
The fixture holds seeded inputs and the reference's bit patterns for the static-wall geometry (SURVEY §8 rows a10, a14,
a16), so the CPU suite can pin O-canon's wall functions without the reference being present."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oref  # noqa: E402


def inputs(seed=11, k=8000):
    """records {rect ox, oy, w, h; a; b; r}.  a = agent centre / ray origin / segment start; b = velocity / ray
    direction / segment end.  Walls are the reference's square tiles (wall_scale 0.3 and 4.0, demo.cpp:164,180) plus
    arbitrary rectangles; the tail of the set is hand-placed edge cases."""
    g = np.random.default_rng(seed)
    rec = np.zeros((k, 9), np.float32)
    side = g.choice(np.array([0.3, 4.0, 1.0], np.float32), k)
    free = g.random(k) < 0.3
    rec[:, 0:2] = g.uniform(-6, 6, (k, 2))
    rec[:, 2] = np.where(free, g.uniform(0.1, 5, k), side)
    rec[:, 3] = np.where(free, g.uniform(0.1, 5, k), side)
    reach = np.maximum(rec[:, 2:4], 1.0)
    rec[:, 4:6] = rec[:, 0:2] + g.uniform(-1.6, 1.6, (k, 2)) * reach      # around and inside the wall
    kind = g.integers(0, 4, k)
    vel = g.uniform(-1, 1, (k, 2))
    vel /= np.maximum(1.0, np.linalg.norm(vel, axis=1, keepdims=True))      # |v| <= 1 (physics.cpp:44-47)
    toward = rec[:, 0:2] - rec[:, 4:6]
    toward = toward / np.maximum(np.linalg.norm(toward, axis=1, keepdims=True), 1e-6) * g.uniform(0.1, 1, (k, 1))
    endpoint = rec[:, 4:6] + g.uniform(-8, 8, (k, 2))
    rec[:, 6:8] = np.where((kind == 0)[:, None], vel, np.where((kind == 1)[:, None], toward,
                           np.where((kind == 2)[:, None], endpoint, vel * g.choice([0.0, 1e-8, 1.0], (k, 2)))))
    rec[:, 8] = np.where(g.random(k) < 0.85, np.float32(0.15), g.uniform(0.05, 0.6, k))
    # agents in contact with a face, a corner, and exactly on the wall centre (normalize(0) -> NaN in the reference)
    m = k - 600
    face = rec[m:m + 200]
    face[:, 4] = face[:, 0] + (face[:, 2] / 2 + face[:, 8] * g.uniform(0.2, 1.1, 200)) * g.choice([-1, 1], 200)
    face[:, 5] = face[:, 1] + face[:, 3] * g.uniform(-0.5, 0.5, 200)
    corner = rec[m + 200:m + 400]
    corner[:, 4] = corner[:, 0] + (corner[:, 2] / 2 + corner[:, 8] * g.uniform(0, 0.8, 200)) * g.choice([-1, 1], 200)
    corner[:, 5] = corner[:, 1] + (corner[:, 3] / 2 + corner[:, 8] * g.uniform(0, 0.8, 200)) * g.choice([-1, 1], 200)
    rec[m + 400:m + 450, 4:6] = rec[m + 400:m + 450, 0:2]
    # axis-parallel rays and segments grazing an edge
    graze = rec[m + 450:]
    graze[:, 4] = graze[:, 0] - graze[:, 2] / 2 - 1.0
    graze[:, 5] = graze[:, 1] + graze[:, 3] / 2 * g.choice([-1, 1, 0.999, 1.001], 150)
    graze[:, 6] = g.choice([1.0, 0.5, -1.0], 150)
    graze[:, 7] = g.choice([0.0, 1e-8, -1e-8], 150)
    return rec


def main():
    rec = inputs()
    out = _oref.run_geom(rec)
    path = os.path.join(ROOT, "tests", "golden", "wall_geometry.npz")
    np.savez_compressed(path, records=rec, words=out)
    f = out.view(np.float32)
    print("records", len(rec), "| ray hits", int((f[:, 0] < 3e38).sum()), "| rect LoS blocked", int((out[:, 1] == 0).sum()),
          "| disc LoS blocked", int((out[:, 2] == 0).sum()), "| ttc <= 5", int((f[:, 3] <= 5).sum()),
          "| contacts", int(out[:, 4].sum()), "| cspace blocked", int((out[:, 7] == 0).sum()),
          "| wall forces", int(out[:, 8].sum()), "| NaN normals", int(np.isnan(f[:, 5]).sum()))


if __name__ == "__main__":
    main()
