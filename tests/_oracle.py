"""ctypes binding of oracle/libcanon.so (O-canon).  TEST INFRASTRUCTURE ONLY — never imported by the product."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
_LIB = None


class Rect(C.Structure):
    _fields_ = [("ox", C.c_float), ("oy", C.c_float), ("w", C.c_float), ("h", C.c_float)]


PLAN_FN = C.CFUNCTYPE(C.c_int32, C.c_void_p, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int32, C.POINTER(C.c_float))
MAX_PLAN = 64  # CANON_MAX_PLAN
BVH_NODE = np.dtype([("ox", np.float32), ("oy", np.float32), ("w", np.float32), ("h", np.float32), ("left", np.int32), ("right", np.int32)])


class Desc(C.Structure):
    _fields_ = [("n", C.c_int32), ("pos", C.c_void_p), ("goal", C.c_void_p), ("radius", C.c_float),
                ("max_duration", C.c_float), ("rng_state", C.c_uint32), ("m_cv", C.c_void_p),
                ("m_f", C.c_void_p), ("comm_enabled", C.c_int32), ("use_grid", C.c_int32),
                ("n_walls", C.c_int32), ("walls", C.c_void_p), ("table_order", C.c_void_p), ("wall_order", C.c_void_p),
                ("plan_fn", PLAN_FN), ("plan_user", C.c_void_p), ("n_bvh", C.c_int32), ("bvh", C.c_void_p), ("features", C.c_uint32)]


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_ORACLE_DIR, "libcanon.so")
        src = os.path.join(_ORACLE_DIR, "canon.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", _ORACLE_DIR, "libcanon.so"], stdout=subprocess.DEVNULL)
        L = C.CDLL(so)
        L.canon_create.restype = C.c_void_p
        L.canon_create.argtypes = [C.POINTER(Desc)]
        L.canon_destroy.argtypes = [C.c_void_p]
        L.canon_step.argtypes = [C.c_void_p, C.c_int32]
        L.canon_step.restype = C.c_int32
        L.canon_step_fixed.argtypes = [C.c_void_p, C.c_int32]
        L.canon_finished.argtypes = [C.c_void_p]
        L.canon_frame.argtypes = [C.c_void_p]
        L.canon_frame.restype = C.c_uint32
        L.canon_get_state.argtypes = [C.c_void_p] + [C.c_void_p] * 8
        L.canon_set_state.argtypes = [C.c_void_p] + [C.c_void_p] * 8
        L.canon_get_rng.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
        L.canon_set_rng.argtypes = [C.c_void_p, C.c_uint32]
        L.canon_get_scores.argtypes = [C.c_void_p] + [C.POINTER(C.c_float)] * 4 + [C.POINTER(C.c_int32)]
        L.canon_get_partners.argtypes = [C.c_void_p, C.c_void_p]
        L.canon_get_neighbor_counts.argtypes = [C.c_void_p, C.c_void_p]
        L.canon_get_neighbor_lists.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.canon_get_neighbor_lists.restype = C.c_int64
        L.canon_overlap_pairs.argtypes = [C.c_void_p]
        L.canon_overlap_pairs.restype = C.c_int64
        L.canon_collision_hits.argtypes = [C.c_void_p]
        L.canon_collision_hits.restype = C.c_int64
        L.canon_record_neighbors.argtypes = [C.c_void_p, C.c_int32]
        L.canon_set_order.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
        L.canon_lcg_next.argtypes = [C.c_uint32]
        L.canon_lcg_next.restype = C.c_uint32
        L.canon_lcg_seed.argtypes = [C.c_uint64]
        L.canon_lcg_seed.restype = C.c_uint32
        L.canon_uniform.argtypes = [C.c_uint32, C.c_float, C.c_float]
        L.canon_uniform.restype = C.c_float
        L.canon_scene.argtypes = [C.c_uint32, C.POINTER(C.c_uint32), C.c_void_p, C.c_void_p,
                                  C.POINTER(C.c_float), C.c_int32]
        L.canon_scene.restype = C.c_int32
        L.canon_scene_full.argtypes = [C.c_uint32, C.POINTER(C.c_uint32), C.c_void_p, C.c_void_p, C.POINTER(C.c_float), C.c_int32,
                                       C.c_void_p, C.POINTER(C.c_int32), C.c_int32]
        L.canon_scene_full.restype = C.c_int32
        L.canon_get_plan.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
        L.canon_get_plan.restype = C.c_int32
        L.canon_get_walled.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.canon_read_config.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p]
        L.canon_write_result.argtypes = [C.c_char_p, C.c_float, C.c_float]
        L.canon_wall_geometry_batch.argtypes = [C.c_int32, C.c_void_p, C.c_void_p]
        for fn in ("canon_rect_intersect", "canon_ttc_circ_wall"):
            getattr(L, fn).restype = C.c_float
        L.canon_rect_intersect.argtypes = [C.POINTER(Rect), C.c_void_p, C.c_void_p]
        L.canon_rect_line_of_sight.argtypes = [C.POINTER(Rect), C.c_void_p, C.c_void_p]
        L.canon_circ_line_of_sight.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]
        L.canon_ttc_circ_wall.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.POINTER(Rect)]
        L.canon_wall_force.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.POINTER(Rect), C.c_void_p]
        L.canon_circ_wall_hit.argtypes = [C.c_void_p, C.c_float, C.POINTER(Rect), C.c_void_p]
        L.canon_cspace_line_of_sight.argtypes = [C.c_void_p, C.c_int32, C.c_float, C.c_void_p, C.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def lookahead(walls, radius, recs):
    """canon_lookahead on _oref.LOOK_REC records -> _oref.LOOK_OUT records."""
    import _oref
    L = lib()
    L.canon_lookahead.argtypes = [C.c_void_p, C.c_int32, C.c_float, C.c_void_p, C.c_void_p, C.c_int32,
                                  C.POINTER(C.c_int32), C.c_void_p]
    walls = np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
    out = np.zeros(len(recs), _oref.LOOK_OUT)
    for k, r in enumerate(recs):
        pos, plan, lg = r["pos"].copy(), np.ascontiguousarray(r["plan"]), r["lg"].copy()
        nd = C.c_int32(int(r["num_done"]))
        L.canon_lookahead(_p(walls), len(walls), radius, _p(pos), _p(plan), int(r["n_plan"]), C.byref(nd), _p(lg))
        out[k] = (nd.value, lg)
    return out


def comm_vis(vis_f, c):
    L = lib()
    L.canon_comm_vis.argtypes = [C.c_int32, C.c_void_p, C.c_void_p]
    cc = np.ascontiguousarray(c, np.float32)
    out = np.zeros(3, np.float32)
    L.canon_comm_vis(vis_f, _p(cc), _p(out))
    return out


def render(pos, c, walls, vis_f=0, cam_dist=1.0, width=500, height=500):
    """canon_render: CPU restatement of render::draw's picture -> uint8 [height][width][4]."""
    L = lib()
    L.canon_render.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_float, C.c_int32, C.c_int32, C.c_void_p]
    pos = np.ascontiguousarray(pos, np.float32)
    c = np.ascontiguousarray(c, np.float32)
    walls = np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
    img = np.zeros((height, width, 4), np.uint8)
    L.canon_render(len(pos), _p(pos), _p(c), len(walls), _p(walls) if len(walls) else None, vis_f, cam_dist, width, height, _p(img))
    return img


def wall_geometry(records):
    """O-canon's static-wall functions on (K, 9) records {rect; a; b; r} -> (K, 12) uint32 words, in the layout of
    oracle/ref/geom_main.cpp."""
    rec = np.ascontiguousarray(records, np.float32)
    out = np.zeros((len(rec), 12), np.uint32)
    lib().canon_wall_geometry_batch(len(rec), _p(rec), _p(out))
    return out


def scene(scene_id, seed=0):
    L = lib()
    rng = C.c_uint32(L.canon_lcg_seed(seed))
    pos = np.zeros((4096, 2), np.float32)
    goal = np.zeros((4096, 2), np.float32)
    md = C.c_float(0)
    n = L.canon_scene(scene_id, C.byref(rng), _p(pos), _p(goal), C.byref(md), 4096)
    assert n > 0, f"scene {scene_id} unsupported by the oracle"
    return pos[:n].copy(), goal[:n].copy(), float(md.value), int(rng.value)


def scene_full(scene_id, seed=0):
    """The whole catalogue incl. walls -> pos, goal, max_duration, rng_state, walls [W][4]."""
    L = lib()
    rng = C.c_uint32(L.canon_lcg_seed(seed))
    pos = np.zeros((4096, 2), np.float32)
    goal = np.zeros((4096, 2), np.float32)
    walls = np.zeros((64, 4), np.float32)
    md, nw = C.c_float(0), C.c_int32(0)
    n = L.canon_scene_full(scene_id, C.byref(rng), _p(pos), _p(goal), C.byref(md), 4096, _p(walls), C.byref(nw), 64)
    assert n > 0, f"scene {scene_id} unsupported by the oracle"
    return pos[:n].copy(), goal[:n].copy(), float(md.value), int(rng.value), walls[:nw.value].copy()


class CanonSim:
    def __init__(self, pos, goal, m_cv=None, m_f=None, max_duration=45.0, rng_state=1, radius=0.15,
                 comm=True, grid=False, walls=None, order=None, wall_order=None, planner=None, bvh=None, features=0):
        """walls [W][4] + planner (an object with plan(start, goal) -> (waypoints [k][2], min_time), e.g. the host planner
        learn_comms_b200.Planner that tests/test_planner.py pins against the reference's own): walled rollouts.
        order / wall_order: table iteration orders (quirk Q4); bvh: flattened static BVH (BVH_NODE records)."""
        L = lib()
        self.n = int(pos.shape[0])
        pos = np.ascontiguousarray(pos, np.float32)
        goal = np.ascontiguousarray(goal, np.float32)
        m_cv = np.zeros((3, 8), np.float32) if m_cv is None else np.ascontiguousarray(m_cv, np.float32)
        m_f = np.zeros((2, 3), np.float32) if m_f is None else np.ascontiguousarray(m_f, np.float32)
        walls = np.zeros((0, 4), np.float32) if walls is None else np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
        order = None if order is None else np.ascontiguousarray(order, np.int32)
        wall_order = None if wall_order is None else np.ascontiguousarray(wall_order, np.int32)
        bvh = None if bvh is None else np.ascontiguousarray(bvh, BVH_NODE)
        self.planner, self.plan_calls = planner, 0

        def _plan(user, agent, start, goal_, cap, out):
            wp, _ = self.planner.plan([start[0], start[1]], [goal_[0], goal_[1]])
            self.plan_calls += 1
            if len(wp) > cap:
                return -1
            for k in range(len(wp)):
                out[2 * k], out[2 * k + 1] = wp[k][0], wp[k][1]
            return len(wp)

        self._cb = PLAN_FN(_plan) if planner is not None else PLAN_FN()
        d = Desc(self.n, _p(pos), _p(goal), radius, max_duration, rng_state, _p(m_cv), _p(m_f),
                 1 if comm else 0, 1 if grid else 0, len(walls), _p(walls) if len(walls) else None, _p(order), _p(wall_order),
                 self._cb, None, 0 if bvh is None else len(bvh), _p(bvh), features)
        self.h = L.canon_create(C.byref(d))
        assert self.h, "canon_create refused the description (walled scene without a planner?)"

    def plan(self, agent):
        wp = np.zeros((MAX_PLAN, 2), np.float32)
        k = lib().canon_get_plan(self.h, agent, _p(wp), MAX_PLAN)
        return wp[:k].copy()

    def walled(self):
        n_plan, lost = np.zeros(self.n, np.int32), np.zeros(self.n, np.int32)
        rp, wh = C.c_int64(0), C.c_int64(0)
        lib().canon_get_walled(self.h, _p(n_plan), _p(lost), C.byref(rp), C.byref(wh))
        return dict(n_plan=n_plan, lost_frames=lost, replans=int(rp.value), wall_hits=int(wh.value))

    def __del__(self):
        if getattr(self, "h", None):
            lib().canon_destroy(self.h)
            self.h = None

    def step(self, n=1, fixed=False):
        return (lib().canon_step_fixed if fixed else lib().canon_step)(self.h, n)

    @property
    def finished(self):
        return bool(lib().canon_finished(self.h))

    @property
    def frame(self):
        return int(lib().canon_frame(self.h))

    def state(self):
        n = self.n
        out = dict(pos=np.zeros((n, 2), np.float32), vel=np.zeros((n, 2), np.float32),
                   vfh=np.zeros((n, 2), np.float32), c=np.zeros((n, 3), np.float32),
                   facing=np.zeros((n, 2), np.float32), local_goal=np.zeros((n, 2), np.float32),
                   goal_dist=np.zeros(n, np.float32), num_done=np.zeros(n, np.int32))
        lib().canon_get_state(self.h, *[_p(out[k]) for k in
                                        ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")])
        return out

    def set_state(self, **kw):
        keys = ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")
        arrs = []
        for k in keys:
            a = kw.get(k)
            if a is not None:
                a = np.ascontiguousarray(a, np.int32 if k == "num_done" else np.float32)
            arrs.append(a)
        lib().canon_set_state(self.h, *[_p(a) for a in arrs])

    @property
    def rng(self):
        v = C.c_uint32(0)
        lib().canon_get_rng(self.h, C.byref(v))
        return int(v.value)

    @rng.setter
    def rng(self, v):
        lib().canon_set_rng(self.h, v)

    def scores(self):
        a, g, u, s = C.c_float(), C.c_float(), C.c_float(), C.c_float()
        k = C.c_int32()
        lib().canon_get_scores(self.h, C.byref(a), C.byref(g), C.byref(u), C.byref(s), C.byref(k))
        return dict(avg_velocity=a.value, group_time=g.value, run_avg=u.value, run_std=s.value, n_done=k.value)

    def partners(self):
        p = np.zeros(self.n, np.int32)
        lib().canon_get_partners(self.h, _p(p))
        return p

    def neighbor_counts(self):
        p = np.zeros(self.n, np.int32)
        lib().canon_get_neighbor_counts(self.h, _p(p))
        return p

    def set_order(self, order=None, gauss_seidel=False):
        o = None if order is None else np.ascontiguousarray(order, np.int32)
        assert lib().canon_set_order(self.h, _p(o), 1 if gauss_seidel else 0) == 0

    def record_neighbors(self, on=True):
        lib().canon_record_neighbors(self.h, 1 if on else 0)

    def neighbor_lists(self):
        tot = lib().canon_get_neighbor_lists(self.h, None, None)
        ptr = np.zeros(self.n + 1, np.int64)
        idx = np.zeros(max(tot, 1), np.int32)
        lib().canon_get_neighbor_lists(self.h, _p(ptr), _p(idx))
        return ptr, idx[:tot]

    @property
    def overlap_pairs(self):
        return int(lib().canon_overlap_pairs(self.h))

    @property
    def collision_hits(self):
        return int(lib().canon_collision_hits(self.h))
