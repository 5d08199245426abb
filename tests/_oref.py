"""Runner / dump reader for oracle/_ref/oref (the reference's own translation units, see oracle/ref/).
TEST INFRASTRUCTURE ONLY.  /root/reference is only needed to BUILD oref; the binary travels with the snapshot."""
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OREF = os.path.join(ROOT, "oracle", "_ref", "oref")
OREF_NOCOMM = os.path.join(ROOT, "oracle", "_ref", "oref_nocomm")


OREF_GEOM = os.path.join(ROOT, "oracle", "_ref", "oref_geom")


def available():
    return os.path.exists(OREF)


def geom_available():
    return os.path.exists(OREF_GEOM)


OREF_PLAN = os.path.join(ROOT, "oracle", "_ref", "oref_plan")


def plan_available():
    return os.path.exists(OREF_PLAN)


def run_plan(walls, radius, start, goal):
    """oracle/ref/plan_main.cpp: the reference's c-space + roadmap + A* -> (vertices, directed edges, [(min_time, waypoints)])."""
    import struct
    walls = np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
    start = np.ascontiguousarray(np.asarray(start, np.float32).reshape(-1, 2))
    goal = np.ascontiguousarray(np.asarray(goal, np.float32).reshape(-1, 2))
    blob = (np.int32(len(walls)).tobytes() + walls.tobytes() + np.float32(radius).tobytes()
            + np.int32(len(start)).tobytes() + start.tobytes() + goal.tobytes())
    out = subprocess.run([OREF_PLAN], input=blob, stdout=subprocess.PIPE, check=True).stdout
    nv, ne = struct.unpack_from("ii", out, 0)
    off, plans = 8, []
    for _ in range(len(start)):
        n, mt = struct.unpack_from("if", out, off)
        off += 8
        plans.append((mt, np.frombuffer(out, np.float32, 2 * n, off).reshape(n, 2).copy()))
        off += 8 * n
    return nv, ne, plans


OREF_LOOK = os.path.join(ROOT, "oracle", "_ref", "oref_look")
LOOK_REC = np.dtype([("pos", np.float32, 2), ("lg", np.float32, 2), ("num_done", np.int32), ("n_plan", np.int32),
                     ("plan", np.float32, (8, 2))])
LOOK_OUT = np.dtype([("num_done", np.int32), ("lg", np.float32, 2)])


def look_available():
    return os.path.exists(OREF_LOOK)


def run_look(walls, radius, recs):
    """oracle/ref/look_main.cpp: the reference's LMP::lookahead on LOOK_REC records -> LOOK_OUT records."""
    walls = np.ascontiguousarray(np.asarray(walls, np.float32).reshape(-1, 4))
    recs = np.ascontiguousarray(recs, LOOK_REC)
    blob = (np.int32(len(walls)).tobytes() + walls.tobytes() + np.float32(radius).tobytes()
            + np.int32(len(recs)).tobytes() + recs.tobytes())
    out = subprocess.run([OREF_LOOK], input=blob, stdout=subprocess.PIPE, check=True).stdout
    return np.frombuffer(out, LOOK_OUT).copy()


OREF_BVH = os.path.join(ROOT, "oracle", "_ref", "oref_bvh")
BVH_NODE = np.dtype([("ox", np.float32), ("oy", np.float32), ("w", np.float32), ("h", np.float32), ("left", np.int32), ("right", np.int32)])


def bvh_available():
    return os.path.exists(OREF_BVH)


def run_bvh(walls_in_statics_order):
    """oracle/ref/bvh_main.cpp: the reference's BVH over the walls (given in statics order) -> pre-order BVH_NODE records;
    leaf.left = position of the wall in the input list."""
    w = np.ascontiguousarray(np.asarray(walls_in_statics_order, np.float32).reshape(-1, 4))
    out = subprocess.run([OREF_BVH], input=np.int32(len(w)).tobytes() + w.tobytes(), stdout=subprocess.PIPE, check=True).stdout
    m = int(np.frombuffer(out, np.int32, 1)[0])
    return np.frombuffer(out, BVH_NODE, m, 4).copy()


def run_geom(records):
    """oracle/ref/geom_main.cpp: (K, 9) float32 records -> (K, 12) uint32 words of the reference's wall geometry."""
    rec = np.ascontiguousarray(records, np.float32)
    blob = np.int32(len(rec)).tobytes() + rec.tobytes()
    out = subprocess.run([OREF_GEOM], input=blob, stdout=subprocess.PIPE, check=True).stdout
    return np.frombuffer(out, np.uint32).reshape(len(rec), 12).copy()


def write_config(path, m_cv, m_f):
    """pso_opt.py:17-23 write_config: 5 lines x 8 floats."""
    cfg = np.zeros((5, 8), np.float64)
    if m_cv is not None:
        cfg[:3] = np.asarray(m_cv, np.float32)
    if m_f is not None:
        cfg[3:, :3] = np.asarray(m_f, np.float32)
    with open(path, "w") as f:
        for row in cfg:
            f.write(" ".join(repr(float(np.float32(v))) for v in row) + "\n")


OREF_NOPROX_NOSOFTSIGN = os.path.join(ROOT, "oracle", "_ref", "oref_noprox_nosoftsign")


def run(scene, seed=0, m_cv=None, m_f=None, max_frames=100000, dump_every=1, comm=True, keep_dir=None, pos_override=None, exe=None):
    """One run of the reference's own translation units.  pos_override: float32 [n][2] start positions that replace the
    scene's own (goals keep the scene's values) — the 1-ulp-perturbation ensembles."""
    import shutil
    d = keep_dir or tempfile.mkdtemp(prefix="oref_")
    write_config(os.path.join(d, "comms.config"), m_cv, m_f)
    dump = os.path.join(d, "dump.bin")
    exe = exe or (OREF if comm else OREF_NOCOMM)
    env = dict(os.environ)
    if pos_override is not None:
        pf = os.path.join(d, "pos_override.f32")
        np.ascontiguousarray(pos_override, np.float32).tofile(pf)
        env["OREF_POS_FILE"] = pf
    subprocess.run([exe, str(scene), str(seed), d, str(max_frames), str(dump_every), dump], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, cwd=d, env=env)
    out = read_dump(dump)
    res = open(os.path.join(d, "comms.result")).read().split()
    out["result_file"] = (res[0], res[1])
    if keep_dir is None:
        shutil.rmtree(d, ignore_errors=True)
    return out


def perturb_ulp(pos, k):
    """Ensemble member k of the 1-ulp perturbation: every coordinate moves by -1, 0 or +1 ulp (member 0 = unperturbed)."""
    pos = np.ascontiguousarray(pos, np.float32)
    if k == 0:
        return pos.copy()
    step = np.random.RandomState(1000 + k).randint(-1, 2, pos.shape)
    out = pos.copy()
    up = np.nextafter(pos, np.float32(np.inf), dtype=np.float32)
    dn = np.nextafter(pos, np.float32(-np.inf), dtype=np.float32)
    out[step > 0] = up[step > 0]
    out[step < 0] = dn[step < 0]
    return out


def read_dump(path):
    buf = np.fromfile(path, np.uint8)
    off = 0

    def take(dtype, count):
        nonlocal off
        nb = np.dtype(dtype).itemsize * count
        a = buf[off:off + nb].view(dtype).copy()
        off += nb
        return a

    magic, n, n_walls, _ = take(np.int32, 4)
    assert magic == 0x4f524546
    out = dict(n=int(n), n_walls=int(n_walls))
    out["order"] = take(np.int32, n) - 1  # entity ids 1..n -> agent indices
    out["pos0"] = take(np.float32, 2 * n).reshape(n, 2)
    out["goal"] = take(np.float32, 2 * n).reshape(n, 2)
    out["min_time"] = take(np.float32, n)
    out["wall_order"] = take(np.int32, n_walls)
    out["walls"] = take(np.float32, 4 * n_walls).reshape(n_walls, 4)
    frames = []
    while True:
        f = int(take(np.int32, 1)[0])
        if f < 0:
            break
        fr = dict(frame=f)
        for k, w in (("pos", 2), ("vel", 2), ("vfh", 2), ("c", 3), ("facing", 2), ("local_goal", 2)):
            fr[k] = take(np.float32, w * n).reshape(n, w)
        fr["goal_dist"] = take(np.float32, n)
        fr["num_done"] = take(np.int32, n)
        fr["run_avg"], fr["run_std"], fr["avg_velocity"] = [float(x) for x in take(np.float32, 3)]
        fr["total_num_done"], fr["all_done"] = [int(x) for x in take(np.int32, 2)]
        frames.append(fr)
    out["frames"] = frames
    out["frames_executed"] = int(take(np.int32, 1)[0])
    av, gt, ra, rs = [float(x) for x in take(np.float32, 4)]
    out["scores"] = dict(avg_velocity=av, group_time=gt, run_avg=ra, run_std=rs, n_done=int(take(np.int32, 1)[0]))
    return out
