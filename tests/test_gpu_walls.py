"""GPU parity of the static-wall geometry (learn_comms_b200/csrc/walls.cuh through lc_wall_geometry) — SURVEY §8 rows
a10 / a14 / a16, first piece of the walled scenes.  Bar: every word equals the reference's own functions
(tests/golden/wall_geometry.npz, made from oracle/_ref/oref_geom) and O-canon; integer words and float bit patterns
exact, a NaN equal to a NaN (x86 and sm_100 produce different quiet-NaN payloads for 0 * inf)."""
import importlib.util
import os

import numpy as np
import pytest

import _oracle
import learn_comms_b200 as lcb

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FLOAT_WORDS = (0, 3, 5, 6, 9, 10)
NAMES = ("Rect::intersect", "Rect::line_of_sight", "Circ::line_of_sight", "LMP::ttc(Circ,Rect)", "circ_rect_collider",
         "normal.x", "normal.y", "Cspace2d::line_of_sight", "wall interacts", "force.x", "force.y", "pad")


def _assert_words(mine, ref):
    for col, name in enumerate(NAMES):
        a, b = mine[:, col], ref[:, col]
        same = a == b
        if col in FLOAT_WORDS:
            same |= np.isnan(a.view(np.float32)) & np.isnan(b.view(np.float32))
        bad = np.nonzero(~same)[0]
        assert bad.size == 0, "%s differs in %d records, first %d" % (name, bad.size, bad[0])


def _generator():
    spec = importlib.util.spec_from_file_location("make_golden_walls", os.path.join(ROOT, "tools", "make_golden_walls.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_wall_geometry_matches_reference_golden():
    g = np.load(os.path.join(ROOT, "tests", "golden", "wall_geometry.npz"))
    _assert_words(lcb.wall_geometry(g["records"]), g["words"])


def test_wall_geometry_matches_oracle_on_fresh_records():
    """1 M records (more than one grid-stride pass of the 148 x 8 blocks) the fixture does not hold."""
    rec = _generator().inputs(seed=77, k=1_000_000)
    _assert_words(lcb.wall_geometry(rec), _oracle.wall_geometry(rec))


def test_wall_geometry_edges_and_errors():
    assert lcb.wall_geometry(np.zeros((0, 9), np.float32)).shape == (0, 12)
    one = np.array([[0, 0, 1, 1, -2, 0, 1, 0, 0.15]], np.float32)  # hand-derived: t = 1.35 s (tests/test_oracle_cpu.py)
    w = lcb.wall_geometry(one)
    assert abs(w.view(np.float32)[0, 3] - 1.35) < 1e-6 and w[0, 8] == 1
    with pytest.raises(lcb.LcError):
        lcb.wall_geometry(np.zeros((4, 8), np.float32))
    with pytest.raises(lcb.LcError):
        lcb.wall_geometry(one, device=99)


def test_lookahead_matches_reference_golden():
    """lc_lookahead (k_lookahead, walls.cu) against the reference's LMP::lookahead on the committed records
    (tests/golden/planner.npz, from oracle/_ref/oref_look): num_done and local_goal, every bit."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "planner.npz"))
    total = 0
    for s in g["seeds"]:
        rin, rout = g[f"look_in{s}"], g[f"look_out{s}"]
        got = lcb.lookahead(g[f"walls{s}"], rin)
        assert got.tobytes() == rout.tobytes(), "layout %d" % s
        total += len(rin)
    assert total > 2000


def test_lookahead_errors_and_no_walls():
    rec = np.zeros(2, lcb.api.LOOK_RECORD)
    rec["n_plan"] = [1, 2]
    rec["plan"][:, 0] = [[5, 0], [0.5, 0]]
    rec["plan"][1, 1] = [9, 9]
    out = lcb.lookahead(np.zeros((0, 4), np.float32), rec)
    assert out["num_done"].tolist() == [0, 1]           # far from its only waypoint; within 1 m of the first, second visible
    assert out["local_goal"].tolist() == [[5, 0], [9, 9]]
    rec["n_plan"][0] = 0
    with pytest.raises(lcb.LcError):
        lcb.lookahead(np.zeros((0, 4), np.float32), rec)
