"""ctypes binding of liblearncomms.so.  Mirrors the reference's system interface for the hot path:
Sim.step() == one iteration of enginee/src/main.cpp:98-150; Sim.scores() == comm::terminate's two numbers."""
import ctypes as C
import os
import numpy as np

from . import build as _build

FLAG_NO_COMM, FLAG_FREE_RUN, FLAG_DEBUG_LISTS = 1, 2, 4
FEAT_PLAIN, FEAT_NO_PROX, FEAT_NO_SOFTSIGN = 32, 64, 128  # comm::x feature-set variants (LC_FLAG_FEAT_*)
_LIB = None


class LcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"learn_comms error {code}: {msg}")
        self.code = code


class _Desc(C.Structure):
    _fields_ = [("n", C.c_int32), ("pos", C.c_void_p), ("goal", C.c_void_p), ("radius", C.c_float),
                ("max_duration", C.c_float), ("rng_state", C.c_uint32), ("m_cv", C.c_void_p), ("m_f", C.c_void_p),
                ("flags", C.c_uint32), ("device", C.c_int32), ("n_walls", C.c_int32), ("walls", C.c_void_p),
                ("stencil", C.c_int32), ("table_order", C.c_void_p), ("wall_order", C.c_void_p)]


class _DomainDesc(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("x_bounds", C.c_void_p), ("capacity", C.c_int32),
                ("halo_capacity", C.c_int32), ("vel", C.c_void_p), ("vel_forehalf", C.c_void_p)]


class _BatchDesc(C.Structure):
    _fields_ = [("n_scenarios", C.c_int32), ("n_agents", C.c_int32), ("pos", C.c_void_p), ("goal", C.c_void_p),
                ("m_cv", C.c_void_p), ("m_f", C.c_void_p), ("max_duration", C.c_void_p), ("rng_state", C.c_void_p),
                ("radius", C.c_float), ("flags", C.c_uint32), ("device", C.c_int32), ("table_order", C.c_void_p),
                ("n_walls", C.c_int32), ("walls", C.c_void_p), ("wall_order", C.c_void_p)]


EXPORTS = [
    "lc_create", "lc_destroy", "lc_set_weights", "lc_step", "lc_run_to_completion", "lc_sync", "lc_finished",
    "lc_get_state", "lc_set_state", "lc_exchange_state", "lc_set_rng", "lc_get_rng", "lc_get_scores", "lc_get_partners",
    "lc_get_neighbor_counts", "lc_get_neighbor_lists", "lc_get_counters", "lc_get_resorts", "lc_timer_start", "lc_timer_stop",
    "lc_kernel_launches", "lc_step_profiled", "lc_count_pair_tests", "lc_fp32_peak", "lc_get_trace", "lc_batch_create", "lc_batch_destroy", "lc_batch_run", "lc_batch_get_state",
    "lc_batch_timer", "lc_batch_reset", "lc_batch_finished", "lc_batch_get_diag", "lc_batch_get_walled", "lc_batch_kernel_launches",
    "lc_dd_create", "lc_dd_strips", "lc_dd_mailbox", "lc_dd_connect", "lc_dd_ipc_handle", "lc_dd_connect_ipc", "lc_dd_get_owned", "lc_dd_counts", "lc_dd_get_partials", "lc_dd_get_shard",
    "lc_read_config", "lc_read_config_legacy", "lc_write_result", "lc_scene", "lc_scene_full", "lc_table_order", "lc_static_bvh", "lc_get_finish_log", "lc_get_walled", "lc_get_plan", "lc_render", "lc_render_static", "lc_wall_geometry", "lc_lookahead", "lc_planner_create", "lc_planner_destroy", "lc_planner_clone",
    "lc_planner_roadmap_size", "lc_planner_plan", "lc_last_error", "lc_version",
]


def lib():
    """Load (building if stale) the CUDA library.  Raises if it cannot be built/loaded: no fallback."""
    global _LIB
    if _LIB is None:
        path = os.environ.get("LEARN_COMMS_LIB") or _build.build()  # override: A/B runs of differently built libraries
        L = C.CDLL(path)
        vp, i32, u32, f32p = C.c_void_p, C.c_int32, C.c_uint32, C.POINTER(C.c_float)
        L.lc_create.argtypes = [C.POINTER(_Desc), C.POINTER(vp)]
        L.lc_destroy.argtypes = [vp]
        L.lc_destroy.restype = None
        L.lc_set_weights.argtypes = [vp, vp, vp]
        L.lc_step.argtypes = [vp, i32]
        L.lc_run_to_completion.argtypes = [vp, C.POINTER(i32)]
        L.lc_sync.argtypes = [vp]
        L.lc_finished.argtypes = [vp, C.POINTER(i32), C.POINTER(u32)]
        L.lc_get_state.argtypes = [vp] + [vp] * 8
        L.lc_set_state.argtypes = [vp] + [vp] * 8
        L.lc_exchange_state.argtypes = [vp] + [vp] * 16
        L.lc_set_rng.argtypes = [vp, u32]
        L.lc_get_rng.argtypes = [vp, C.POINTER(u32)]
        L.lc_get_scores.argtypes = [vp, f32p, f32p, f32p, f32p, C.POINTER(i32)]
        L.lc_get_partners.argtypes = [vp, vp]
        L.lc_get_neighbor_counts.argtypes = [vp, vp]
        L.lc_get_neighbor_lists.argtypes = [vp, vp, i32]
        L.lc_dd_strips.argtypes = [i32, vp, i32, i32, C.c_float, vp]
        L.lc_get_counters.argtypes = [vp] + [C.POINTER(C.c_int64)] * 3
        L.lc_get_resorts.argtypes = [vp, C.POINTER(C.c_int64)]
        L.lc_timer_start.argtypes = [vp]
        L.lc_timer_stop.argtypes = [vp, f32p]
        L.lc_kernel_launches.argtypes = [vp]
        L.lc_kernel_launches.restype = C.c_int64
        L.lc_step_profiled.argtypes = [vp, i32, vp]
        L.lc_count_pair_tests.argtypes = [vp, C.POINTER(C.c_int64)]
        L.lc_get_trace.argtypes = [vp, vp]
        L.lc_fp32_peak.argtypes = [i32, i32, vp]
        L.lc_dd_create.argtypes = [C.POINTER(_Desc), C.POINTER(_DomainDesc), C.POINTER(vp)]
        L.lc_dd_mailbox.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_uint64)]
        L.lc_dd_connect.argtypes = [vp, i32, vp, i32]
        L.lc_dd_ipc_handle.argtypes = [vp, vp]
        L.lc_dd_connect_ipc.argtypes = [vp, i32, vp]
        L.lc_dd_get_owned.argtypes = [vp, vp]
        L.lc_dd_counts.argtypes = [vp, C.POINTER(i32), C.POINTER(i32)]
        L.lc_dd_get_shard.argtypes = [vp, vp, vp, vp, i32, C.POINTER(i32)]
        L.lc_dd_get_partials.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(i32), f32p]
        L.lc_batch_create.argtypes = [C.POINTER(_BatchDesc), C.POINTER(vp)]
        L.lc_batch_destroy.argtypes = [vp]
        L.lc_batch_destroy.restype = None
        L.lc_batch_run.argtypes = [vp, i32, vp, vp, vp, vp]
        L.lc_batch_get_state.argtypes = [vp] + [vp] * 8
        L.lc_batch_timer.argtypes = [vp, f32p]
        L.lc_batch_reset.argtypes = [vp, vp, vp]
        L.lc_batch_finished.argtypes = [vp, C.POINTER(i32)]
        L.lc_batch_get_diag.argtypes = [vp] + [vp] * 5
        L.lc_batch_get_walled.argtypes = [vp, vp, vp, vp, C.POINTER(C.c_int64)]
        L.lc_batch_kernel_launches.argtypes = [vp]
        L.lc_batch_kernel_launches.restype = C.c_int64
        L.lc_read_config.argtypes = [C.c_char_p, vp, vp]
        L.lc_read_config_legacy.argtypes = [C.c_char_p, vp, vp, C.POINTER(u32)]
        L.lc_write_result.argtypes = [C.c_char_p, C.c_float, C.c_float]
        L.lc_scene.argtypes = [u32, C.c_uint64, vp, vp, i32, f32p, C.POINTER(u32)]
        L.lc_scene_full.argtypes = [u32, C.c_uint64, vp, vp, i32, f32p, C.POINTER(u32), vp, i32, C.POINTER(i32)]
        L.lc_get_finish_log.argtypes = [vp, vp, i32, C.POINTER(i32)]
        L.lc_get_walled.argtypes = [vp, vp, vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.lc_get_plan.argtypes = [vp, i32, vp, i32, C.POINTER(i32)]
        L.lc_render.argtypes = [vp, i32, C.c_float, i32, i32, vp]
        L.lc_render_static.argtypes = [vp, i32, C.c_float, i32, i32, vp]
        L.lc_table_order.argtypes = [i32, i32, vp, vp]
        L.lc_static_bvh.argtypes = [i32, vp, vp, vp, i32, C.POINTER(i32)]
        L.lc_wall_geometry.argtypes = [i32, vp, vp, i32]
        L.lc_lookahead.argtypes = [i32, vp, C.c_float, i32, vp, vp, i32]
        L.lc_planner_create.argtypes = [i32, vp, C.c_float, C.POINTER(vp)]
        L.lc_planner_clone.argtypes = [vp, C.POINTER(vp)]
        L.lc_planner_destroy.argtypes = [vp]
        L.lc_planner_destroy.restype = None
        L.lc_planner_roadmap_size.argtypes = [vp, C.POINTER(i32), C.POINTER(i32)]
        L.lc_planner_plan.argtypes = [vp, vp, vp, i32, vp, C.POINTER(i32), f32p]
        L.lc_last_error.restype = C.c_char_p
        L.lc_version.restype = C.c_char_p
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _ck(rc):
    if rc != 0:
        raise LcError(rc, lib().lc_last_error().decode())


def scene(scene_id, seed=0):
    """demo::init's scene catalogue (demo.cpp:141-511), wall-free scenes.  Returns pos, goal, max_duration, rng_state."""
    pos = np.zeros((4096, 2), np.float32)
    goal = np.zeros((4096, 2), np.float32)
    md, rng = C.c_float(0), C.c_uint32(0)
    n = lib().lc_scene(scene_id, seed, _p(pos), _p(goal), 4096, C.byref(md), C.byref(rng))
    if n <= 0:
        raise LcError(n, f"scene {scene_id} is not in the wall-free catalogue")
    return pos[:n].copy(), goal[:n].copy(), float(md.value), int(rng.value)


def dd_strips(pos, world, ghost_band):
    """lc_dd_strips: the x cuts lc_dd_create picks by default (local counts — owned + ghosts — equalised).  Host code."""
    pos = np.ascontiguousarray(pos, np.float32)
    xb = np.zeros(world + 1, np.float32)
    _ck(lib().lc_dd_strips(pos.shape[0], _p(pos), 2, world, C.c_float(ghost_band), _p(xb)))
    return xb


def scene_full(scene_id, seed=0):
    """The whole catalogue incl. walled scenes (lc_scene_full) -> pos, goal, max_duration, rng_state, walls [W][4]."""
    pos = np.zeros((4096, 2), np.float32)
    goal = np.zeros((4096, 2), np.float32)
    walls = np.zeros((64, 4), np.float32)
    md, rng, nw = C.c_float(0), C.c_uint32(0), C.c_int32(0)
    n = lib().lc_scene_full(scene_id, seed, _p(pos), _p(goal), 4096, C.byref(md), C.byref(rng), _p(walls), 64, C.byref(nw))
    if n <= 0:
        raise LcError(n, f"scene {scene_id} is not in the catalogue")
    return pos[:n].copy(), goal[:n].copy(), float(md.value), int(rng.value), walls[:nw.value].copy()


def table_order(n_agents, n_walls=0):
    """Iteration order of the reference's component tables (quirk Q4, lc_table_order): (agent_order, wall_order)."""
    a = np.zeros(max(n_agents, 1), np.int32)
    w = np.zeros(max(n_walls, 1), np.int32)
    _ck(lib().lc_table_order(n_agents, n_walls, _p(a), _p(w)))
    return a[:n_agents].copy(), w[:n_walls].copy()


BVH_NODE = np.dtype([("ox", np.float32), ("oy", np.float32), ("w", np.float32), ("h", np.float32), ("left", np.int32), ("right", np.int32)])


def static_bvh(walls, wall_order=None):
    """lc_static_bvh: the static BVH of ai::init over the walls (statics in wall_order), pre-order BVH_NODE records."""
    w = np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
    o = None if wall_order is None else np.ascontiguousarray(wall_order, np.int32)
    out = np.zeros(2 * len(w), BVH_NODE)
    m = C.c_int32(0)
    _ck(lib().lc_static_bvh(len(w), _p(w), _p(o), _p(out), len(out), C.byref(m)))
    return out[:m.value].copy()


def _order(order, n):
    if order is None:
        return None
    if isinstance(order, str):
        assert order == "reference"
        return table_order(n)[0]
    o = np.ascontiguousarray(order, np.int32)
    assert o.shape == (n,)
    return o


def wall_geometry(records, device=0):
    """lc_wall_geometry: (K, 9) float32 records {wall ox, oy, w, h; a; b; r} -> (K, 12) uint32 words, computed on the
    device (BoundVolume.cpp:76-230, LMP.cpp:64-82,307-319, BVH.cpp:121-130, PRM.cpp:133-173)."""
    rec = np.ascontiguousarray(records, np.float32)
    if rec.ndim != 2 or rec.shape[1] != 9:
        raise LcError(-1, "wall_geometry: records must be (K, 9)")
    out = np.zeros((len(rec), 12), np.uint32)
    _ck(lib().lc_wall_geometry(len(rec), _p(rec), _p(out), device))
    return out


LOOK_RECORD = np.dtype([("pos", np.float32, 2), ("local_goal", np.float32, 2), ("num_done", np.int32), ("n_plan", np.int32),
                        ("plan", np.float32, (8, 2))])   # lc_look_record
LOOK_RESULT = np.dtype([("num_done", np.int32), ("local_goal", np.float32, 2)])  # lc_look_result


def lookahead(walls, records, agent_radius=0.15, device=0):
    """lc_lookahead: LMP::lookahead (LMP.cpp:84-113) on the device for LOOK_RECORD records -> LOOK_RESULT records."""
    w = np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
    rec = np.ascontiguousarray(records)
    if rec.dtype.itemsize != LOOK_RECORD.itemsize:
        raise LcError(-1, "lookahead: records must have the lc_look_record layout")
    out = np.zeros(len(rec), LOOK_RESULT)
    _ck(lib().lc_lookahead(len(w), _p(w), agent_radius, len(rec), _p(rec), _p(out), device))
    return out


class Planner:
    """Host-side global planner for walled scenes (lc_planner_*): Cspace2d + PRM roadmap + A* as ai::init builds them
    (ai.cpp:41-63, PRM.cpp:8-105, GMP.cpp:7-110).  Stateful like the reference's shared roadmap: call plan() in the
    reference's order."""

    def __init__(self, walls, agent_radius=0.15):
        self._w = np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
        self._h = C.c_void_p()
        _ck(lib().lc_planner_create(len(self._w), _p(self._w), agent_radius, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().lc_planner_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def clone(self):
        """Independent planner with the same roadmap state (lc_planner_clone)."""
        c = Planner.__new__(Planner)
        c._w = self._w
        c._h = C.c_void_p()
        _ck(lib().lc_planner_clone(self._h, C.byref(c._h)))
        return c

    def roadmap_size(self):
        v, e = C.c_int32(0), C.c_int32(0)
        _ck(lib().lc_planner_roadmap_size(self._h, C.byref(v), C.byref(e)))
        return v.value, e.value

    def plan(self, start, goal, cap=256):
        """-> (waypoints [n][2], min_time); n = 0 when the goal is unreachable."""
        s, g = np.asarray(start, np.float32).reshape(2).copy(), np.asarray(goal, np.float32).reshape(2).copy()
        wp = np.zeros((cap, 2), np.float32)
        n, mt = C.c_int32(0), C.c_float(0)
        _ck(lib().lc_planner_plan(self._h, _p(s), _p(g), cap, _p(wp), C.byref(n), C.byref(mt)))
        return wp[:n.value].copy(), float(mt.value)


def read_config(path):
    m_cv, m_f = np.zeros((3, 8), np.float32), np.zeros((2, 3), np.float32)
    rc = lib().lc_read_config(os.fsencode(path), _p(m_cv), _p(m_f))
    if rc:
        raise LcError(rc, f"cannot parse {path} (comm.cpp:80-121 format: 5 lines x 8 floats)")
    return m_cv, m_f


def read_config_legacy(path):
    """lc_read_config_legacy: a "3 5"-headed legacy file -> (m_cv [3][8], m_f [2][3], feature flags)."""
    m_cv, m_f = np.zeros((3, 8), np.float32), np.zeros((2, 3), np.float32)
    fl = C.c_uint32(0)
    rc = lib().lc_read_config_legacy(os.fsencode(path), _p(m_cv), _p(m_f), C.byref(fl))
    if rc:
        raise LcError(rc, f"cannot parse {path} as a legacy 3 x 5 config (optimizer.py:20-27)")
    return m_cv, m_f, int(fl.value)


def write_result(path, avg_velocity, group_time):
    rc = lib().lc_write_result(os.fsencode(path), avg_velocity, group_time)
    if rc:
        raise LcError(rc, f"cannot write {path}")


_STATE_KEYS = ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")


class Sim:
    """One crowd resident on one GPU.  step(k) runs k frames of main.cpp:98-150."""

    def __init__(self, pos, goal, m_cv=None, m_f=None, max_duration=45.0, rng_state=1, radius=0.15, comm=True,
                 free_run=False, debug_lists=False, device=0, stencil=0, order=None, walls=None, wall_order=None, features=0):
        """order: table iteration order (lc_desc::table_order) — None = ascending index, "reference" = the reference's
        own (lc_table_order), or an explicit permutation.  walls [W][4] (origin x, z, width, height): a walled scene;
        wall_order = statics order (None = creation order; with order="reference" the reference's own)."""
        self.n = int(pos.shape[0])
        self._walls = None if walls is None or len(walls) == 0 else np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
        nw = 0 if self._walls is None else len(self._walls)
        if isinstance(order, str) and nw and wall_order is None:
            wall_order = table_order(self.n, nw)[1]
        self._wall_order = None if (wall_order is None or nw == 0) else np.ascontiguousarray(wall_order, np.int32)
        self._order = _order(order, self.n)
        pos = np.ascontiguousarray(pos, np.float32)
        goal = np.ascontiguousarray(goal, np.float32)
        m_cv = None if m_cv is None else np.ascontiguousarray(m_cv, np.float32)
        m_f = None if m_f is None else np.ascontiguousarray(m_f, np.float32)
        flags = (0 if comm else FLAG_NO_COMM) | (FLAG_FREE_RUN if free_run else 0) | (FLAG_DEBUG_LISTS if debug_lists else 0) | features
        d = _Desc(self.n, _p(pos), _p(goal), radius, max_duration, rng_state, _p(m_cv), _p(m_f), flags, device, nw,
                  _p(self._walls), stencil, _p(self._order), _p(self._wall_order))
        h = C.c_void_p()
        self.h = None
        _ck(lib().lc_create(C.byref(d), C.byref(h)))
        self.h = h
        self.list_cap = min(self.n - 1, 256) if debug_lists else 0

    def close(self):
        if getattr(self, "h", None):
            lib().lc_destroy(self.h)
            self.h = None

    __del__ = close

    def set_weights(self, m_cv, m_f):
        m_cv = np.ascontiguousarray(m_cv, np.float32)
        m_f = np.ascontiguousarray(m_f, np.float32)
        _ck(lib().lc_set_weights(self.h, _p(m_cv), _p(m_f)))

    def step(self, n=1, sync=True):
        _ck(lib().lc_step(self.h, n))
        if sync:
            _ck(lib().lc_sync(self.h))

    def sync(self):
        _ck(lib().lc_sync(self.h))

    def run_to_completion(self):
        k = C.c_int32(0)
        _ck(lib().lc_run_to_completion(self.h, C.byref(k)))
        return int(k.value)

    @property
    def finished(self):
        f, k = C.c_int32(0), C.c_uint32(0)
        _ck(lib().lc_finished(self.h, C.byref(f), C.byref(k)))
        return bool(f.value)

    @property
    def frame(self):
        f, k = C.c_int32(0), C.c_uint32(0)
        _ck(lib().lc_finished(self.h, C.byref(f), C.byref(k)))
        return int(k.value)

    def state(self, keys=_STATE_KEYS, out=None):
        n = self.n
        shapes = dict(pos=(n, 2), vel=(n, 2), vfh=(n, 2), c=(n, 3), facing=(n, 2), local_goal=(n, 2), goal_dist=(n,),
                      num_done=(n,))
        out = {} if out is None else out
        for k in keys:
            if k not in out:
                out[k] = np.zeros(shapes[k], np.int32 if k == "num_done" else np.float32)
        _ck(lib().lc_get_state(self.h, *[_p(out.get(k)) if k in keys else None for k in _STATE_KEYS]))
        return out

    def set_state(self, **kw):
        arrs = []
        for k in _STATE_KEYS:
            a = kw.get(k)
            if a is not None:
                a = np.ascontiguousarray(a, np.int32 if k == "num_done" else np.float32)
            arrs.append(a)
        _ck(lib().lc_set_state(self.h, *[_p(a) for a in arrs]))
        _ck(lib().lc_sync(self.h))

    def exchange_state(self, out, **kw):
        """state() into the preallocated arrays of `out` and set_state(**kw), as one call whose two copies overlap
        (lc_exchange_state).  Keys missing from `out` / kw are skipped."""
        ins = []
        for k in _STATE_KEYS:
            a = kw.get(k)
            if a is not None:
                a = np.ascontiguousarray(a, np.int32 if k == "num_done" else np.float32)
            ins.append(a)
        _ck(lib().lc_exchange_state(self.h, *[_p(out[k]) if k in out else None for k in _STATE_KEYS], *[_p(a) for a in ins]))
        return out

    @property
    def rng(self):
        v = C.c_uint32(0)
        _ck(lib().lc_get_rng(self.h, C.byref(v)))
        return int(v.value)

    @rng.setter
    def rng(self, v):
        _ck(lib().lc_set_rng(self.h, v))

    def scores(self):
        a, g, u, s = C.c_float(), C.c_float(), C.c_float(), C.c_float()
        k = C.c_int32()
        _ck(lib().lc_get_scores(self.h, C.byref(a), C.byref(g), C.byref(u), C.byref(s), C.byref(k)))
        return dict(avg_velocity=a.value, group_time=g.value, run_avg=u.value, run_std=s.value, n_done=k.value)

    def partners(self):
        p = np.zeros(self.n, np.int32)
        _ck(lib().lc_get_partners(self.h, _p(p)))
        return p

    def neighbor_counts(self):
        p = np.zeros(self.n, np.int32)
        _ck(lib().lc_get_neighbor_counts(self.h, _p(p)))
        return p

    def neighbor_lists(self):
        """CSR (row_ptr, idx) of {j : ttc_ij <= 5} for the last frame; needs debug_lists=True."""
        cnt = self.neighbor_counts()
        assert cnt.max(initial=0) <= self.list_cap, "neighbour list capacity exceeded"
        dense = np.zeros((self.n, self.list_cap), np.int32)
        _ck(lib().lc_get_neighbor_lists(self.h, _p(dense), self.list_cap))
        ptr = np.zeros(self.n + 1, np.int64)
        np.cumsum(cnt, out=ptr[1:])
        idx = np.concatenate([dense[i, :cnt[i]] for i in range(self.n)]) if self.n else np.zeros(0, np.int32)
        return ptr, idx.astype(np.int32)  # rows by agent index; entries in table order (ascending index by default)

    def counters(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        _ck(lib().lc_get_counters(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(overlap_pairs=a.value, collision_hits=b.value, candidates=c.value)

    def resorts(self):
        """Frames so far that re-sorted the grid (the others kept slots — and, decomposed, owners and ghosts)."""
        a = C.c_int64()
        _ck(lib().lc_get_resorts(self.h, C.byref(a)))
        return a.value

    def timer_start(self):
        _ck(lib().lc_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        _ck(lib().lc_timer_stop(self.h, C.byref(ms)))
        return float(ms.value)

    KERNELS = ("k_frame_begin", "k_bin", "k_scan_hull", "k_reorder", "k_refine", "k_pairs", "k_force", "k_collide")

    def step_profiled(self, n=1):
        """n frames launched kernel by kernel with CUDA events in between -> {kernel: mean ms per launch}."""
        ms = np.zeros(len(self.KERNELS), np.float32)
        _ck(lib().lc_step_profiled(self.h, n, _p(ms)))
        return dict(zip(self.KERNELS, ms.tolist()))

    def render(self, vis_f=0, cam_dist=1.0, width=500, height=500):
        """lc_render: RGBA8 picture [height][width][4] of the crowd coloured by comm_vis(c) (render.cpp:87-172,364-415)."""
        img = np.zeros((height, width, 4), np.uint8)
        _ck(lib().lc_render(self.h, vis_f, cam_dist, width, height, _p(img)))
        return img

    def render_static(self, vis_f=0, cam_dist=1.0, width=500, height=500, out=None):
        """lc_render_static: the picture of render() through the entry point that allocates nothing on the device (static
        scratch, at most 2^20 pixels).  `out` = optional preallocated uint8 array [height][width][4]."""
        img = np.zeros((height, width, 4), np.uint8) if out is None else out
        _ck(lib().lc_render_static(self.h, vis_f, cam_dist, width, height, _p(img)))
        return img

    def finish_log(self):
        """[k][3] (overhead, running mean, running sigma) of every counted finisher, in Welford order (demo.cpp:655-663)."""
        log = np.zeros((self.n, 3), np.float32)
        k = C.c_int32(0)
        _ck(lib().lc_get_finish_log(self.h, _p(log), self.n, C.byref(k)))
        return log[:k.value].copy()

    def walled(self):
        n_plan, lost = np.zeros(self.n, np.int32), np.zeros(self.n, np.int32)
        rp, wh = C.c_int64(0), C.c_int64(0)
        _ck(lib().lc_get_walled(self.h, _p(n_plan), _p(lost), C.byref(rp), C.byref(wh)))
        return dict(n_plan=n_plan, lost_frames=lost, replans=int(rp.value), wall_hits=int(wh.value))

    def plan(self, agent):
        wp = np.zeros((64, 2), np.float32)
        k = C.c_int32(0)
        _ck(lib().lc_get_plan(self.h, agent, _p(wp), 64, C.byref(k)))
        return wp[:k.value].copy()

    def trace(self):
        """Device-clock stamps (ns) of the last executed frame (lc_get_trace): dict name -> ns relative to the frame's start."""
        ts = np.zeros(16, np.uint64)
        _ck(lib().lc_get_trace(self.h, _p(ts)))
        names = ("begin", "bin", "scan_hull", "reorder", "refine", "pairs", "force", "collide", "nbr_wait_begin", "nbr_wait_end",
                 "all_wait_begin", "all_wait_end", "refine_end", "begin_end")
        t0 = int(ts[0])
        return {k: (int(ts[i]) - t0 if ts[i] else None) for i, k in enumerate(names)}

    def pair_tests(self):
        """pair-filter evaluations of the last executed frame (lc_count_pair_tests)."""
        k = C.c_int64(0)
        _ck(lib().lc_count_pair_tests(self.h, C.byref(k)))
        return int(k.value)

    @property
    def launches(self):
        return int(lib().lc_kernel_launches(self.h))


def fp32_peak(device=0, repeats=5):
    """Measured FP32 TFLOP/s (lc_fp32_peak): dict(mul_add = scalar uncontracted FMUL + FADD chains, packed = the same as
    f32x2 instructions, fma = FFMA chains)."""
    t = np.zeros(3, np.float32)
    _ck(lib().lc_fp32_peak(device, repeats, _p(t)))
    return dict(mul_add=float(t[0]), packed=float(t[1]), fma=float(t[2]))


class DomainSim(Sim):
    """One rank (x strip) of a crowd that is spatially decomposed over several GPUs (lc_dd_*).  Every rank is built from
    the same global arrays.  Ranks of one process are wired with connect_local(); one process per GPU (torchrun) with
    connect_dist().  state() returns global-size arrays that are zero outside the rank's own agents plus `owned`."""

    def __init__(self, pos, goal, rank, world, m_cv=None, m_f=None, max_duration=45.0, rng_state=1, radius=0.15, comm=True,
                 free_run=False, device=0, stencil=0, x_bounds=None, capacity=0, halo_capacity=0, vel=None, vfh=None, order=None):
        self.n = int(pos.shape[0])
        self._order = _order(order, self.n)
        self.rank, self.world, self.device = int(rank), int(world), int(device)
        pos = np.ascontiguousarray(pos, np.float32)
        goal = np.ascontiguousarray(goal, np.float32)
        m_cv = None if m_cv is None else np.ascontiguousarray(m_cv, np.float32)
        m_f = None if m_f is None else np.ascontiguousarray(m_f, np.float32)
        xb = None if x_bounds is None else np.ascontiguousarray(x_bounds, np.float32)
        vel = None if vel is None else np.ascontiguousarray(vel, np.float32)
        vfh = None if vfh is None else np.ascontiguousarray(vfh, np.float32)
        flags = (0 if comm else FLAG_NO_COMM) | (FLAG_FREE_RUN if free_run else 0)
        d = _Desc(self.n, _p(pos), _p(goal), radius, max_duration, rng_state, _p(m_cv), _p(m_f), flags, device, 0, None, stencil,
                  _p(self._order), None)
        dom = _DomainDesc(rank, world, _p(xb), capacity, halo_capacity, _p(vel), _p(vfh))
        h = C.c_void_p()
        self.h = None
        _ck(lib().lc_dd_create(C.byref(d), C.byref(dom), C.byref(h)))
        self.h = h
        self.list_cap = 0

    def mailbox(self):
        ptr, nbytes = C.c_void_p(), C.c_uint64()
        _ck(lib().lc_dd_mailbox(self.h, C.byref(ptr), C.byref(nbytes)))
        return ptr, int(nbytes.value)

    def owned(self):
        o = np.zeros(self.n, np.int32)
        _ck(lib().lc_dd_get_owned(self.h, _p(o)))
        return o

    def counts(self):
        a, b = C.c_int32(), C.c_int32()
        _ck(lib().lc_dd_counts(self.h, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def set_state(self, **kw):
        raise LcError(-1, "set_state is not available for a domain-decomposed crowd")

    def shard(self, out=None):
        """(ids, pos, vel) of the agents this rank owns right now, compact; `out` = preallocated (pinned) arrays."""
        cap = self.counts()[1] if out is None else out["ids"].shape[0]
        if out is None:
            out = dict(ids=np.zeros(cap, np.int32), pos=np.zeros((cap, 2), np.float32), vel=np.zeros((cap, 2), np.float32))
        k = C.c_int32(0)
        _ck(lib().lc_dd_get_shard(self.h, _p(out["ids"]), _p(out["pos"]), _p(out["vel"]), cap, C.byref(k)))
        return out, int(k.value)

    def partials(self):
        lv, nf, av = C.c_uint64(), C.c_int32(), C.c_float()
        _ck(lib().lc_dd_get_partials(self.h, C.byref(lv), C.byref(nf), C.byref(av)))
        return int(lv.value), int(nf.value), float(av.value)

    @staticmethod
    def merge_scores(rank0_scores, partials):
        """comm::terminate's numbers of the whole crowd: Welford statistics from the scoring rank, avg_velocity
        (physics.cpp:72-76: the speed of the last not-done agent in table order) from the ranks' partials."""
        out = dict(rank0_scores)
        key = max(p[0] for p in partials)
        if key:
            sp = np.array([key & 0xffffffff], np.uint32).view(np.float32)[0]
            others = sum(p[1] for p in partials) - (0 if np.isfinite(sp) else 1)
            poisoned = others > 0 or not np.isfinite(partials[0][2])
            out["avg_velocity"] = float("nan") if poisoned else float(sp)
        return out

    @staticmethod
    def merged_scores(sims):
        return DomainSim.merge_scores(sims[0].scores(), [s.partials() for s in sims])

    def scores_dist(self, dist):
        """One process per GPU: every rank returns the whole crowd's scores."""
        import torch
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        sc = self.scores()
        lv, nf, av = self.partials()
        keys = sorted(sc)
        mine = torch.tensor([float(lv >> 32), float(lv & 0xffffffff), float(nf)] + [float(sc[k]) for k in keys],
                            dtype=torch.float64, device=dev)
        parts = [torch.zeros_like(mine) for _ in range(self.world)]
        dist.all_gather(parts, mine)
        parts = [p.cpu().numpy() for p in parts]
        r0 = {k: (int(parts[0][3 + i]) if k == "n_done" else float(np.float32(parts[0][3 + i]))) for i, k in enumerate(keys)}
        partials = [((int(p[0]) << 32) | int(p[1]), int(p[2]), float(np.float32(parts[0][3 + keys.index("avg_velocity")]))) for p in parts]
        return DomainSim.merge_scores(r0, partials)

    @staticmethod
    def connect_local(sims):
        """Ranks living in this process: hand every rank the device pointer of every other rank's mailbox."""
        for a in sims:
            for b in sims:
                if a is not b:
                    ptr, _ = b.mailbox()
                    _ck(lib().lc_dd_connect(a.h, b.rank, ptr, b.device))

    def connect_dist(self, dist):
        """One process per GPU: all-gather the 64-byte CUDA IPC handles of the mailboxes over torch.distributed."""
        import torch
        mine = np.zeros(64, np.uint8)
        _ck(lib().lc_dd_ipc_handle(self.h, _p(mine)))
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        t = torch.from_numpy(mine).to(dev)
        parts = [torch.zeros_like(t) for _ in range(self.world)]
        dist.all_gather(parts, t)
        for r in range(self.world):
            if r != self.rank:
                hnd = np.ascontiguousarray(parts[r].cpu().numpy())
                _ck(lib().lc_dd_connect_ipc(self.h, r, _p(hnd)))
        dist.barrier()

    @staticmethod
    def step_all(sims, frames, chunk=16):
        """Same-process ranks: enqueue `frames` frames on every rank, interleaved in chunks so that no stream's launch
        queue fills up while its kernels wait for a peer whose work has not been enqueued yet."""
        left = frames
        while left > 0:
            k = min(chunk, left)
            for sm in sims:
                sm.step(k, sync=False)
            left -= k
            for sm in sims:
                sm.sync()

    @staticmethod
    def merged_state(sims):
        """Global tables assembled from the owners' entries (same-process ranks)."""
        out, seen = None, None
        for sm in sims:
            st = sm.state()
            own = sm.owned().astype(bool)
            if out is None:
                out = {k: np.zeros_like(v) for k, v in st.items()}
                seen = np.zeros(sm.n, np.int32)
            seen += own
            for k, v in st.items():
                out[k][own] = v[own]
        assert (seen == 1).all(), f"ownership is not a partition: {np.bincount(seen)}"
        return out


class Batch:
    """Many independent small scenarios in one launch (pso_opt.py:236-267 simulate_set, in process)."""

    def __init__(self, pos, goal, m_cv, m_f, max_duration, rng_state, radius=0.15, comm=True, device=0, order=None, walls=None,
                 wall_order=None, features=0):
        """walls [W][4]: one wall layout shared by every scenario of the batch (replans are served by the host planner
        between relaunches of the rollout kernel); order / wall_order as for Sim."""
        S, n = pos.shape[0], pos.shape[1]
        self.S, self.n = int(S), int(n)
        self._walls = None if walls is None or len(walls) == 0 else np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
        nw = 0 if self._walls is None else len(self._walls)
        if isinstance(order, str) and nw and wall_order is None:
            wall_order = table_order(self.n, nw)[1]
        self._wall_order = None if (wall_order is None or nw == 0) else np.ascontiguousarray(wall_order, np.int32)
        self._order = _order(order, self.n)
        self._keep = [np.ascontiguousarray(pos, np.float32), np.ascontiguousarray(goal, np.float32),
                      np.ascontiguousarray(m_cv, np.float32), np.ascontiguousarray(m_f, np.float32),
                      np.ascontiguousarray(np.broadcast_to(np.asarray(max_duration, np.float32), (S,))),
                      np.ascontiguousarray(np.broadcast_to(np.asarray(rng_state, np.uint32), (S,)))]
        assert self._keep[2].shape == (S, 3, 8) and self._keep[3].shape == (S, 2, 3)
        d = _BatchDesc(S, n, *[_p(a) for a in self._keep], radius, (0 if comm else FLAG_NO_COMM) | features, device, _p(self._order),
                       nw, _p(self._walls), _p(self._wall_order))
        h = C.c_void_p()
        self.h = None
        _ck(lib().lc_batch_create(C.byref(d), C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            lib().lc_batch_destroy(self.h)
            self.h = None

    __del__ = close

    def run(self, max_frames=(1 << 31) - 1):
        """Advance every unfinished scenario by up to max_frames frames (default: each to its own all_done)."""
        S = self.S
        out = dict(avg_velocity=np.zeros(S, np.float32), group_time=np.zeros(S, np.float32),
                   n_done=np.zeros(S, np.int32), frames=np.zeros(S, np.int32))
        _ck(lib().lc_batch_run(self.h, max_frames, _p(out["avg_velocity"]), _p(out["group_time"]), _p(out["n_done"]),
                               _p(out["frames"])))
        return out

    def reset(self, m_cv=None, m_f=None):
        """Frame 0 again, optionally with new per-scenario parameters (the next PSO iteration)."""
        if m_cv is not None:
            m_cv = np.ascontiguousarray(m_cv, np.float32)
            assert m_cv.shape == (self.S, 3, 8)
        if m_f is not None:
            m_f = np.ascontiguousarray(m_f, np.float32)
            assert m_f.shape == (self.S, 2, 3)
        _ck(lib().lc_batch_reset(self.h, _p(m_cv), _p(m_f)))

    @property
    def n_finished(self):
        k = C.c_int32(0)
        _ck(lib().lc_batch_finished(self.h, C.byref(k)))
        return int(k.value)

    def state(self):
        S, n = self.S, self.n
        out = dict(pos=np.zeros((S, n, 2), np.float32), vel=np.zeros((S, n, 2), np.float32),
                   vfh=np.zeros((S, n, 2), np.float32), c=np.zeros((S, n, 3), np.float32),
                   facing=np.zeros((S, n, 2), np.float32), local_goal=np.zeros((S, n, 2), np.float32),
                   goal_dist=np.zeros((S, n), np.float32), num_done=np.zeros((S, n), np.int32))
        _ck(lib().lc_batch_get_state(self.h, *[_p(out[k]) for k in _STATE_KEYS]))
        return out

    def diag(self):
        S, n = self.S, self.n
        out = dict(partner=np.zeros((S, n), np.int32), neighbor_count=np.zeros((S, n), np.int32),
                   overlap_pairs=np.zeros(S, np.int64), collision_hits=np.zeros(S, np.int64),
                   pair_tests=np.zeros(S, np.int64))
        _ck(lib().lc_batch_get_diag(self.h, *[_p(out[k]) for k in ("partner", "neighbor_count", "overlap_pairs",
                                                                     "collision_hits", "pair_tests")]))
        return out

    def walled(self):
        S, n = self.S, self.n
        out = dict(n_plan=np.zeros((S, n), np.int32), lost_frames=np.zeros((S, n), np.int32), wall_hits=np.zeros(S, np.int64))
        rp = C.c_int64(0)
        _ck(lib().lc_batch_get_walled(self.h, _p(out["n_plan"]), _p(out["lost_frames"]), _p(out["wall_hits"]), C.byref(rp)))
        out["replans"] = int(rp.value)
        return out

    def device_ms(self):
        ms = C.c_float()
        _ck(lib().lc_batch_timer(self.h, C.byref(ms)))
        return float(ms.value)

    @property
    def launches(self):
        return int(lib().lc_batch_kernel_launches(self.h))
