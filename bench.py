#!/usr/bin/env python
"""bench.py — agent-steps/sec of the learn-comms per-frame loop on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload plaza1m|corridor100k|circle1k|scene0|pso4096] [--impl reference]

One "step" = one frame (main.cpp:98-150: demo::run, comm::run, ai::update_agents, physics::simulate, all_sync) of the
whole crowd.  Default workload: BASELINE.json configs[4], the 1M-agent open plaza the north_star's target is quoted
on (it fits one GPU).  `value` = agent-steps/s with the state resident in HBM; `e2e` = the same frame driven through
the C ABI with HOST component tables inside the timed region; `e2e_read` = lc_step(1) + host read of (pos, vel), the
one definition that exists at every N.  `roofline` is for the dominant pass (k_pairs + k_force), timed live with CUDA
events on the launch stream, with an FP32 leg (counted pair tests x 22 flop against the measured CUDA-core peak).
`state_hash` = order-independent 64-bit hash of (id, pos, vel) of every agent after exactly warmup + steps frames: the
same value at every N is the bit-identity claim of the domain decomposition, shown on real peers.
`secondary` carries the metric's second half (PSO scenario evals/s) from the pso4096 batch, sharded over the same N GPUs.
`cpu_baseline` / `--impl reference` time the reference's CPU loop on ALL host cores: O-ref (the reference's own
translation units) where it can hold the workload, else O-canon (its CPU restatement; the reference's entity store is
limited to 65 535 entities, Entity.h:12, and its loop is O(N^2)).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly ONE JSON line.  Libraries write there too (NCCL prints its "NCCL version ..." banner on fd 1),
# so fd 1 is pointed at stderr for the whole run and the result line goes to a private duplicate of the real stdout.
sys.stdout.flush()
RESULT_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)

ALG_BYTES_PER_AGENT_STEP = 112  # 60 B compulsory read + 52 B compulsory write of live agent state
FLOP_PER_PAIR_TEST = 22         # 2 vector subtractions, 3 dot products, b*b - a*c, compares
WORKLOADS = {
    # name: (generator kwargs, description, cpu-sample kwargs)
    "plaza1m": dict(kind="plaza", n=1_000_000, side=2000.0, seed=5, wscale=0.1,
                    desc="open plaza 2000x2000 m, 1,000,000 agents (0.25/m^2), random goals, seed 5",
                    sample=dict(n=62_500, side=500.0)),
    "plaza125k": dict(kind="plaza", n=125_000, side=707.1, seed=5, wscale=0.1,
                      desc="1/8 of the plaza at the same density (development: the per-GPU load of the 8-GPU decomposition)",
                      sample=dict(n=62_500, side=500.0)),
    "corridor100k": dict(kind="corridor", cols=500, rows=100, seed=4, wscale=0.1,
                         desc="bidirectional corridor, 2 x (500x100) lattice blocks = 100,000 agents, seed 4",
                         sample=dict(cols=125, rows=100)),
    "pso4096": dict(kind="pso", n_scenarios=4096, n_agents=64, seed=3, wscale=1.0,
                    desc="PSO fitness batch: 4096 independent scenarios (512 particles x 8 wall-free layouts) x 64 agents, "
                         "per-particle comm parameters, each rolled out to its own all_done (<= 45 s = 2701 frames)",
                    sample=dict(n_scenarios=8)),
    "pso512": dict(kind="pso", n_scenarios=512, n_agents=64, seed=3, wscale=1.0,
                   desc="development: the first 512 scenarios of pso4096 = one rank's share of it at 8 GPUs", sample=dict(n_scenarios=8)),
    "circle1k": dict(kind="circle", n=1000, radius=80.0, seed=1, wscale=0.1,
                     desc="circle antipodal swap, 1,000 agents, r = 80 m, seed 1", sample=dict(n=1000, radius=80.0)),
    "scene0": dict(kind="scene", scene=0, seed=0, wscale=1.0,
                   desc="the reference's own scene 0 (demo.cpp:145-151: 60 agents on a 10 m circle, antipodal goals, <= 45 s), "
                        "learned-weight rollouts to completion in the reference's table order", sample=dict()),
}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def make_workload(name, sample=False, seed_offset=0):
    from learn_comms_b200 import workloads as wl
    w = dict(WORKLOADS[name])
    kw = {k: v for k, v in w.items() if k not in ("kind", "desc", "sample", "wscale")}
    if sample:
        kw.update(w["sample"])
    kw["seed"] = kw["seed"] + seed_offset
    pos, goal, md, rng = getattr(wl, w["kind"])(**kw)
    m_cv, m_f = wl.weights(w["seed"], w["wscale"])
    return pos, goal, md, rng, m_cv, m_f


def walking_state(pos, goal):
    """Developed-flow initial condition for both arms: every agent already walks at unit speed towards its goal
    (vel = vel_forehalf = normalize(goal - pos)), so the interaction cutoff 0.3 + 10 v_max is at its steady value
    from the first timed frame instead of after a ~2 s ramp from rest."""
    d = (goal - pos).astype(np.float32)
    l = np.maximum(np.hypot(d[:, 0], d[:, 1]), np.float32(1e-6)).astype(np.float32)
    v = (d / l[:, None]).astype(np.float32)
    return dict(vel=v, vfh=v.copy())


def state_hash_partial(ids, pos, vel):
    """Sum over agents (mod 2^64) of a 64-bit mix of (id, pos bits, vel bits): independent of the order agents are
    listed in and of how they are split over ranks."""
    with np.errstate(over="ignore"):
        h = (ids.astype(np.uint64) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15)
        for k, col in enumerate((pos[:, 0], pos[:, 1], vel[:, 0], vel[:, 1])):
            b = np.ascontiguousarray(col, np.float32).view(np.uint32).astype(np.uint64)
            h ^= (b + np.uint64(k + 1)) * np.uint64(0xBF58476D1CE4E5B9)
            h = (h ^ (h >> np.uint64(29))) * np.uint64(0x94D049BB133111EB)
        h ^= h >> np.uint64(32)
        return int(h.sum(dtype=np.uint64))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,timestamp")

    def __init__(self, index):
        self.index, self.proc, self.path, self.t_mark = index, None, None, None

    def mark(self):
        """Samples taken before this moment are dropped (the sampler may be started early: nvidia-smi is slow to come up)."""
        self.t_mark = time.time()

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        time.sleep(0.15)
        t_end = time.time() + 3.0
        late = False
        while time.time() < t_end and os.path.getsize(self.path) == 0:   # no line yet (slow start with 8 samplers at once): wait for one
            late = True
            time.sleep(0.05)
        if late:
            out["note"] = "first sample arrived after the sampled window ended"
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            ts = None
            if len(f) >= 10:
                try:
                    import datetime
                    ts = datetime.datetime.strptime(f[9], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                except ValueError:
                    ts = None
            rows.append((ts, f))
        if self.t_mark is not None:
            inside = [r for r in rows if r[0] is not None and r[0] >= self.t_mark - 0.02]
            if inside:
                rows = inside
            elif rows:
                rows = rows[-1:]
                out["note"] = "no sample inside the window; the last one before it"
        for ts, f in rows:
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nme, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm),
                       power_w_max=float(max(pw)))
        return out


def dist_setup(n_gpus):
    """torch.distributed plumbing (NCCL) when launched by torchrun; returns (rank, world, local_rank, dist or None)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1:
        return 0, 1, 0, None
    import torch
    import torch.distributed as dist
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local, dist


def hbm_peak():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_entry(name, key):
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        return json.load(open(tp)).get(name, {}).get(key)
    return None


# ---------------------------------------------------------------------------------------------------------------------
# CPU arms (test infrastructure executed ONLY here: the cpu_baseline leg and --impl reference)

def _oracle_mod():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    return _oracle


def crowd_cpu_rate(name, frames, warm, threads):
    """O-canon (uniform-grid search) on the bounded sample of `name`, one independent sample crowd per host thread
    (different seeds), all stepping `frames` frames at once: aggregate agent-steps/s of the box's host cores."""
    _oracle = _oracle_mod()
    brute = WORKLOADS[name]["kind"] == "circle"   # the reference's own all-pairs loop where it is affordable
    sims, n = [], 0
    for t in range(threads):
        pos, goal, md, rng, m_cv, m_f = make_workload(name, sample=True, seed_offset=100 * t)
        o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, grid=not brute)
        o.set_state(**walking_state(pos, goal))
        sims.append(o)
        n = pos.shape[0]
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda o: o.step(warm, fixed=True), sims))
        t0 = time.perf_counter()
        list(ex.map(lambda o: o.step(frames, fixed=True), sims))
        dt = time.perf_counter() - t0
    return threads * n * frames / dt, n, dt


def sample_text(name, threads):
    w = WORKLOADS[name]
    how = "brute force (the reference's all-pairs loop)" if w["kind"] == "circle" else "uniform-grid search"
    return (f"{w['kind']} sample {w['sample']} of the workload at the same density, O-canon {how}, one independent sample crowd "
            f"per host thread x {threads} threads")


def oref_rate(scene, threads, runs_per_thread=1):
    """The reference's own translation units (oracle/_ref/oref) on a catalogue scene, rollouts to completion, one process
    per host thread (pso_opt.py's own fan-out, pso_opt.py:236-242).  Returns (agent-steps/s, evals/s, frames, seconds) or None."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oref
    if not _oref.available():
        return None
    from learn_comms_b200 import workloads as wl
    m_cv, m_f = wl.weights(1)
    jobs = list(range(threads * runs_per_thread))

    def one(k):
        r = _oref.run(scene, k, m_cv, m_f, dump_every=0)
        return r["n"] * r["frames_executed"]

    with ThreadPoolExecutor(threads) as ex:
        t0 = time.perf_counter()
        steps = sum(ex.map(one, jobs))
        dt = time.perf_counter() - t0
    return steps / dt, len(jobs) / dt, steps, dt


def pso_cpu_rate(n_scenarios, n_agents, seed, count, threads):
    """O-canon brute force (the reference's all-pairs loop) on `count` scenarios spread evenly over the batch, one
    rollout per host thread at a time."""
    _oracle = _oracle_mod()
    from learn_comms_b200 import workloads as wl
    pos, goal, m_cv, m_f, md, rng = wl.pso_batch(n_scenarios, n_agents, seed=seed)
    pick = [int(k * n_scenarios / count) for k in range(count)]

    def one(s):
        o = _oracle.CanonSim(pos[s], goal[s], m_cv[s], m_f[s], max_duration=float(md[s]), rng_state=int(rng[s]))
        o.step(10 ** 6)
        return o.frame * n_agents

    with ThreadPoolExecutor(threads) as ex:
        t0 = time.perf_counter()
        steps = sum(ex.map(one, pick))
        dt = time.perf_counter() - t0
    return steps / dt, count / dt, count, dt


def reference_line(args, value, ms_per_step, config, cpu, extra=None):
    line = {"impl": "reference", "metric": "agent_steps_per_sec", "value": value, "unit": "agent-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": cpu, "e2e": {"value": value, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if extra:
        line.update(extra)
    print(json.dumps(line), file=RESULT_OUT, flush=True)


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores, all of them."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    name, T = args.workload, host_cores()
    w = WORKLOADS[name]
    if w["kind"] == "scene":
        # the reference's own binary can hold this workload: O-ref, one process per core
        tot_steps, tot_dt, tot_evals = 0.0, 0.0, 0
        for k in range(args.warmup + args.steps):
            r = oref_rate(w["scene"], T)
            if r is None:
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/oref is not built (needs /root/reference at build time)"}),
                      file=RESULT_OUT, flush=True)
                return
            if k >= args.warmup:
                tot_steps += r[2]; tot_dt += r[3]; tot_evals += T
        v = tot_steps / tot_dt
        cpu = {"value": v, "unit": "agent-steps/s", "cores": T, "kind": "reference",
               "sample": f"oracle/_ref/oref (the reference's unmodified translation units), scene {w['scene']}, seeds 0..{T - 1}, one process per core per step"}
        return reference_line(args, v, 1e3 * tot_dt / args.steps, {"workload": name, "desc": w["desc"]}, cpu,
                              {"scenario_evals_per_sec": tot_evals / tot_dt})
    if w["kind"] == "pso":
        S, n = w["n_scenarios"], w["n_agents"]
        tot_steps, tot_evals, tot_dt = 0.0, 0, 0.0
        for k in range(args.warmup + args.steps):
            v, ev, evals, dt = pso_cpu_rate(S, n, w["seed"], T, T)
            if k >= args.warmup:
                tot_steps += v * dt; tot_evals += evals; tot_dt += dt
        v = tot_steps / tot_dt
        cpu = {"value": v, "unit": "agent-steps/s", "cores": T, "kind": "port",
               "sample": f"{T} of the {S} scenarios per step (evenly spread), O-canon brute force, one rollout per host thread"}
        return reference_line(args, v, 1e3 * tot_dt / args.steps,
                              {"workload": name, "desc": w["desc"], "scenarios": S, "agents_per_scenario": n}, cpu,
                              {"scenario_evals_per_sec": tot_evals / tot_dt})
    v, n, dt = crowd_cpu_rate(name, args.steps, args.warmup, T)
    cpu = {"value": v, "unit": "agent-steps/s", "cores": T, "kind": "port", "sample": sample_text(name, T),
           "why_port": "the reference's entity store holds <= 65 535 entities (Entity.h:12) and its loops are O(N^2): it cannot run this workload"}
    reference_line(args, v, 1e3 * dt / args.steps, {"workload": name, "desc": w["desc"], "agents": n * T, "frames_per_step": 1}, cpu)


# ---------------------------------------------------------------------------------------------------------------------
def batch_measure(name, K, W, rank, world, local, dist, with_e2e=True):
    """PSO fitness batch (BASELINE.json configs[2]): one step = every scenario of the batch rolled out to completion in
    ONE kernel launch per GPU.  N > 1: contiguous blocks of scenarios per rank, no data-path collective, scores gathered
    (16 B per scenario); total work is fixed => strong scaling.  Returns the measurement dict (rank 0 prints it)."""
    import torch
    import learn_comms_b200 as lcb
    from learn_comms_b200 import workloads as wl
    from learn_comms_b200 import pso
    w = WORKLOADS[name]
    S, n = w["n_scenarios"], w["n_agents"]
    pos, goal, m_cv, m_f, md, rng = wl.pso_batch(S, n, seed=w["seed"], scale=w["wscale"])
    lo, hi = pso.shard_range(S, rank, world)
    sl = slice(lo, hi)
    b = lcb.Batch(pos[sl], goal[sl], m_cv[sl], m_f[sl], md[sl], rng[sl], device=local)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(W):
        b.reset()
        out = b.run()
    barrier()
    dev_ms, frames_sum, l0 = 0.0, 0, b.launches
    for _ in range(K):
        b.reset()
        out = b.run()
        dev_ms += b.device_ms()   # CUDA events around the rollout kernel on its own stream
        frames_sum += int(out["frames"].sum())
    barrier()
    launches = b.launches - l0
    pair_tests = int(b.diag()["pair_tests"].sum())
    steps_local = frames_sum * n
    if dist is not None:
        t = torch.tensor([dev_ms], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); dev_ms = float(t.item())
        t = torch.tensor([steps_local, pair_tests, launches], device="cuda", dtype=torch.int64); dist.all_reduce(t)
        steps_all, pair_tests, launches = int(t[0].item()), int(t[1].item()), int(t[2].item())
    else:
        steps_all = steps_local
    value = steps_all / (dev_ms * 1e-3)
    evals = S * K / (dev_ms * 1e-3)
    e2e = None
    if with_e2e:
        # host parameter matrices in, host scores out, through the evaluator pso_opt.py would call (reset = H2D of the
        # initial tables + weights, run = rollouts + D2H of the score block), wall clock, max over ranks
        Ke = max(3, min(K, 10))
        ev = pso.BatchEvaluator(pos, goal, md, rng, device=local, rank=rank, world=world, dist=dist, order=None)
        ev.evaluate(m_cv, m_f)
        barrier()
        t0 = time.perf_counter()
        fr = 0
        for _ in range(Ke):
            sc = ev.evaluate(m_cv, m_f)
            fr += int(sc["frames"].sum())
        barrier()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = float(t.item())
        e2e = {"value": fr * n / dt, "unit": "agent-steps/s", "h2d_bytes_per_step": ev.h2d_bytes * world, "d2h_bytes_per_step": ev.d2h_bytes * world,
               "steps": Ke, "ms_per_step": 1e3 * dt / Ke, "scenario_evals_per_sec": S * Ke / dt,
               "api": "pso.BatchEvaluator.evaluate(host M_cv[S], M_f[S]) -> host scores: lc_batch_reset + lc_batch_run"}
    peak, peak_src = hbm_peak()
    achieved = ALG_BYTES_PER_AGENT_STEP * (steps_all / world) / (dev_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "k_batch_rollout", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic_entry(name, "k_batch_rollout_dram_bytes_per_launch"), "peak_source": peak_src, "kernel_share": 1.0,
                "pair_tests_per_sec": pair_tests * K / (dev_ms * 1e-3),
                "note": "state lives in shared memory / registers for the whole rollout: HBM is touched once per scenario, the kernel "
                        "is FP32-issue bound on the all-pairs TTC loop (DESIGN.md); algorithmic bytes = 112 B x agent-steps"}
    return {"agent_steps_per_sec": value, "scenario_evals_per_sec": evals, "ms_per_step": dev_ms / K, "steps": K, "warmup": W,
            "gpu_launches": int(launches), "roofline": roofline, "e2e": e2e, "scenarios": S, "agents_per_scenario": n, "desc": w["desc"]}


def run_batch(args):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    rank, world, local, dist = dist_setup(args.gpus)
    W, K = max(args.warmup, 3), args.steps
    clocks = ClockSampler(local)
    clocks.start()
    m = batch_measure(args.workload, K, W, rank, world, local, dist)
    clk = clocks.stop()
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        T = host_cores()
        w = WORKLOADS[args.workload]
        v, evs, k, dtc = pso_cpu_rate(m["scenarios"], m["agents_per_scenario"], w["seed"], max(T, 16), T)
        cpu = {"value": v, "unit": "agent-steps/s", "cores": T, "kind": "port", "scenario_evals_per_sec": evs,
               "sample": f"{k} of the {m['scenarios']} scenarios (evenly spread), O-canon brute force, one rollout per host thread, {dtc:.1f} s"}
    if rank == 0:
        line = {"metric": "agent_steps_per_sec", "value": m["agent_steps_per_sec"], "unit": "agent-steps/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "scenario_evals_per_sec": m["scenario_evals_per_sec"],
                "config": {"workload": args.workload, "desc": m["desc"], "scenarios": m["scenarios"], "agents_per_scenario": m["agents_per_scenario"],
                           "parallelism": f"{world} GPU(s), contiguous scenario blocks, score gather only",
                           "l2": "no HBM working set: state is staged once into shared memory; nothing to flush"},
                "clocks": clk, "e2e": m["e2e"], "gpu_launches": m["gpu_launches"], "roofline": m["roofline"], "cpu_baseline": cpu}
        print(json.dumps(line), file=RESULT_OUT, flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def run_scene(args):
    """BASELINE.json configs[0]-style workload: the reference's own scene, rollouts to completion through lc_create /
    lc_run_to_completion in the reference's table order; one step = one rollout."""
    import torch
    import learn_comms_b200 as lcb
    from learn_comms_b200 import workloads as wl
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    rank, world, local, dist = dist_setup(args.gpus)
    w = WORKLOADS[args.workload]
    W, K = max(args.warmup, 3), args.steps
    m_cv, m_f = wl.weights(1)
    steps, ms, launches = 0, 0.0, 0
    clocks = ClockSampler(local)
    clocks.start()
    for k in range(W + K):
        pos, goal, md, rng = lcb.scene(w["scene"], k + 1000 * rank)
        sim = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, device=local, order="reference")
        sim.timer_start()
        frames = sim.run_to_completion()
        t = sim.timer_stop()
        if k >= W:
            steps += frames * pos.shape[0]; ms += t; launches += sim.launches
        sim.close()
    clk = clocks.stop()
    if dist is not None:
        t = torch.tensor([ms], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
        t = torch.tensor([steps, launches], device="cuda", dtype=torch.int64); dist.all_reduce(t)
        steps, launches = int(t[0].item()), int(t[1].item())
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        T = host_cores()
        r = oref_rate(w["scene"], T)
        if r is not None:
            cpu = {"value": r[0], "unit": "agent-steps/s", "cores": T, "kind": "reference", "scenario_evals_per_sec": r[1],
                   "sample": f"oracle/_ref/oref (the reference's unmodified translation units), scene {w['scene']}, seeds 0..{T - 1}, one process per core, {r[3]:.1f} s"}
    if rank == 0:
        peak, peak_src = hbm_peak()
        achieved = ALG_BYTES_PER_AGENT_STEP * (steps / world) / (ms * 1e-3) / 1e9
        line = {"metric": "agent_steps_per_sec", "value": steps / (ms * 1e-3), "unit": "agent-steps/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "scenario_evals_per_sec": K * world / (ms * 1e-3),
                "config": {"workload": args.workload, "desc": w["desc"], "parallelism": f"{world} independent rollouts at a time, one per GPU",
                           "l2": "60 agents: the whole state is cache resident; launch latency bound (the batched kernel is the tool for this size)"},
                "clocks": clk, "e2e": None, "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                             "peak_source": peak_src, "note": "8 launches of latency per frame for 60 agents; nowhere near any roofline"},
                "cpu_baseline": cpu}
        print(json.dumps(line), file=RESULT_OUT, flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default="plaza1m", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--stencil", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the PSO-batch leg of the default line")
    ap.add_argument("--replicas", action="store_true", help="N > 1: N independent crowds instead of one decomposed crowd")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if WORKLOADS[args.workload]["kind"] == "pso":
        return run_batch(args)
    if WORKLOADS[args.workload]["kind"] == "scene":
        return run_scene(args)

    import torch
    import learn_comms_b200 as lcb
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    rank, world, local, dist = dist_setup(args.gpus)
    W, K = max(args.warmup, 3), args.steps
    name = args.workload
    # N > 1: ONE crowd of the named shape, decomposed into N x strips (one per GPU); boundary agents, comm-partner
    # candidates and the global reductions travel through peer memory over NVLink inside the frame kernels (domain.cuh)
    domain = world > 1 and not args.replicas
    clocks = ClockSampler(local)   # started now, counted from clocks.mark() on: with 8 samplers at once nvidia-smi takes > 1 s to come up
    clocks.start()
    pos, goal, md, rng, m_cv, m_f = make_workload(name, seed_offset=0 if domain else rank)
    n = pos.shape[0]
    if domain:
        ws = walking_state(pos, goal)
        sim = lcb.DomainSim(pos, goal, rank, world, m_cv, m_f, max_duration=md, rng_state=rng, free_run=True, device=local,
                            stencil=args.stencil, vel=ws["vel"], vfh=ws["vfh"])
        sim.connect_dist(dist)
    else:
        sim = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, free_run=True, device=local, stencil=args.stencil)
        sim.set_state(**walking_state(pos, goal))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        sim.sync()

    # nvidia-smi needs a few hundred ms for its first line and the timed region can be as short as 2 ms (20 frames on 8 GPUs):
    # the sampler starts before the warm-up steps and runs through the timed region and the frames that follow it (timeline,
    # per-kernel profiling: the same frames of the same crowd), so its samples are taken under this very load
    clocks.mark()
    sim.step(W)
    barrier()
    l0 = sim.launches
    barrier()
    sim.timer_start()
    sim.step(K, sync=False)
    ms = sim.timer_stop()
    barrier()
    launches = sim.launches - l0
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    n_total = n if domain else world * n   # agents advanced per step by the whole job
    value = n_total * K / (ms * 1e-3)
    n_local = sim.counts()[0] if domain else n
    n_owned = int(sim.owned().sum()) if domain else n

    # bit-identity across N: hash of (id, pos, vel) of every agent after exactly W + K frames (the same crowd at every N)
    if domain:
        sh, k_own = sim.shard()
        part = state_hash_partial(sh["ids"][:k_own], sh["pos"][:k_own], sh["vel"][:k_own])
        t = torch.tensor([part >> 32, part & 0xffffffff, k_own], device="cuda", dtype=torch.int64)
        parts = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(parts, t)
        total = sum((int(p[0].item()) << 32) | int(p[1].item()) for p in parts) & ((1 << 64) - 1)
        hashed = sum(int(p[2].item()) for p in parts)
    else:
        st = sim.state(keys=("pos", "vel"))
        total = state_hash_partial(np.arange(n, dtype=np.int64), st["pos"], st["vel"]) & ((1 << 64) - 1)
        hashed = n
    state_hash = {"hash": f"{total:016x}", "frames": W + K, "agents_hashed": hashed,
                  "what": "sum mod 2^64 over all agents of mix64(id, pos bits, vel bits); equal at every N <=> the decomposed run is bit-identical"}

    # timeline of one frame INSIDE the graph replay (device clock stamps at kernel entries; the two peer waits of a decomposed
    # crowd): what the frame's critical path really is — per-kernel CUDA events need a serial, host-paced launch sequence
    sim.step(16)
    tr = sim.trace()
    order = ["begin", "bin", "scan_hull", "reorder", "pairs", "force", "collide"]
    seg = {f"{a}->{b_}": (tr[b_] - tr[a]) / 1e3 for a, b_ in zip(order[:-1], order[1:]) if tr[a] is not None and tr[b_] is not None}
    timeline = {"unit": "us", "entry_to_entry": seg, "refine_entry": None if tr["refine"] is None else tr["refine"] / 1e3,
                "refine_end": None if tr["refine_end"] is None else tr["refine_end"] / 1e3}
    if domain:
        mine = {"rank": rank, "nbr_wait": (tr["nbr_wait_end"] - tr["nbr_wait_begin"]) / 1e3, "all_wait": (tr["all_wait_end"] - tr["all_wait_begin"]) / 1e3,
                "begin_kernel": (tr["begin_end"] - tr["begin"]) / 1e3 if tr["begin_end"] else None, "entry_to_entry": seg,
                "refine_end": timeline["refine_end"], "force_entry": tr["force"] / 1e3}
        allr = [None] * world
        dist.all_gather_object(allr, mine)
        timeline["ranks"] = allr
        timeline["nbr_wait_max"] = max(r["nbr_wait"] for r in allr)
        timeline["all_wait_max"] = max(r["all_wait"] for r in allr)

    # roofline leg: per-kernel device time, CUDA events between the launches on the simulation's stream
    prof = sim.step_profiled(20)
    clk = clocks.stop()
    clk["window"] = "warm-up + timed region + the timeline / per-kernel profiling frames right after it (same crowd, same kernels)"
    pair_tests = sim.pair_tests()
    fp = lcb.fp32_peak(local)
    t_force = (prof["k_pairs"] + prof["k_force"]) * 1e-3  # S1-S3a: pair filter + force/integrate, one logical pass over the state
    peak, peak_src = hbm_peak()
    achieved = ALG_BYTES_PER_AGENT_STEP * n_owned / t_force / 1e9  # this rank's own agents per k_pairs + k_force launch
    frame_ms = sum(prof.values())
    fp_achieved = pair_tests * FLOP_PER_PAIR_TEST / (prof["k_pairs"] * 1e-3) / 1e12
    roofline = {"bound": "hbm", "kernel": "k_pairs+k_force", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic_entry(name, "k_pairs_k_force_dram_bytes_per_launch"), "peak_source": peak_src, "kernel_ms": prof,
                "kernel_share": (prof["k_pairs"] + prof["k_force"]) / frame_ms,
                "frame": {"achieved": ALG_BYTES_PER_AGENT_STEP * n_owned / (ms / K * 1e-3) / 1e9,
                          "frac": ALG_BYTES_PER_AGENT_STEP * n_owned / (ms / K * 1e-3) / 1e9 / peak,
                          "traffic": traffic_entry(name, "frame_dram_bytes"), "algorithmic_bytes": ALG_BYTES_PER_AGENT_STEP * n_owned},
                "fp32": {"kernel": "k_pairs", "pair_tests_per_launch": pair_tests, "flop_per_pair_test": FLOP_PER_PAIR_TEST,
                         "achieved": fp_achieved, "unit": "TFLOP/s",
                         "peak_fma": fp["fma"], "peak_mul_add_scalar": fp["mul_add"], "peak_mul_add_packed": fp["packed"],
                         "frac_of_fma_peak": fp_achieved / fp["fma"] if fp["fma"] else None,
                         "frac_of_packed_peak": fp_achieved / fp["packed"] if fp["packed"] else None,
                         "pair_tests_per_agent": pair_tests / max(n_local, 1),
                         "peak_source": "lc_fp32_peak, measured in this run: independent chains, no memory traffic; the library may not "
                                        "contract mul+add (bit parity with the reference's SSE2 build), so the packed mul/add figure is its ceiling"},
                "note": "pair loop is FP32/issue bound, not HBM bound (DESIGN.md); algorithmic bytes = 112 B x agents per launch"}

    # e2e legs.  e2e_read (every N): lc_step(1) + host read of (pos, vel) [+ ids on a decomposed crowd], pinned host tables.
    # e2e: the host engine owns the component tables — at N = 1 the full round trip lc_set_state -> lc_step(1) -> lc_get_state;
    # teacher forcing is single-GPU only (learn_comms.h), so on a decomposed crowd e2e is e2e_read.
    e2e, e2e_read = None, None
    if not args.no_e2e:
        Ke = max(3, min(K, 50))
        if domain:
            cap = sim.counts()[1]
            pinned = dict(ids=torch.zeros(cap, dtype=torch.int32).pin_memory(), pos=torch.zeros((cap, 2)).pin_memory(),
                          vel=torch.zeros((cap, 2)).pin_memory())
            views = {k: v.numpy() for k, v in pinned.items()}
            read = lambda: sim.shard(out=views)[1] * 20 + 4  # noqa: E731
            api = "lc_step(1) + lc_dd_get_shard (ids, pos, vel of the rank's own agents) per rank"
        else:
            pinned = dict(pos=torch.zeros((n, 2)).pin_memory(), vel=torch.zeros((n, 2)).pin_memory())
            views = {k: v.numpy() for k, v in pinned.items()}
            read = lambda: (sim.state(keys=("pos", "vel"), out=views), n * 16)[1]  # noqa: E731
            api = "lc_step(1) + lc_get_state(pos, vel)"
        for _ in range(3):
            sim.step(1, sync=False)
            nb = read()
        barrier()
        t0 = time.perf_counter()
        for _ in range(Ke):
            sim.step(1, sync=False)
            nb = read()
        barrier()
        dt = time.perf_counter() - t0
        d2h = nb
        if dist is not None:
            t = torch.tensor([dt], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = float(t.item())
            t = torch.tensor([float(nb)], device="cuda"); dist.all_reduce(t); d2h = int(t.item()) if domain else nb * world
        e2e_read = {"value": n_total * Ke / dt, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": int(d2h),
                    "steps": Ke, "ms_per_step": 1e3 * dt / Ke, "api": api + ", pinned host tables; the state stays device-resident"}
        if domain:
            e2e = dict(e2e_read, note="decomposed crowd: lc_set_state (teacher forcing) is single-GPU only, so e2e = e2e_read here; "
                                      "the like-for-like host round trip is the N = 1 line's e2e")
        else:
            st = sim.state()
            pinned_t = {k: torch.from_numpy(a.copy()).pin_memory() for k, a in st.items()}
            pinned = {k: t.numpy() for k, t in pinned_t.items()}
            h2d = sum(a.nbytes for a in pinned.values())
            for _ in range(3):
                sim.set_state(**pinned)
                sim.step(1)
                sim.state(out=pinned)
            barrier()
            t0 = time.perf_counter()
            for _ in range(Ke):
                sim.set_state(**pinned)
                sim.step(1, sync=False)
                sim.state(out=pinned)
            barrier()
            dt = time.perf_counter() - t0
            if dist is not None:
                t = torch.tensor([dt], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = float(t.item())
            e2e_dep = {"value": n_total * Ke / dt, "unit": "agent-steps/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": h2d * world,
                       "steps": Ke, "ms_per_step": 1e3 * dt / Ke, "host_gb_per_s": 2 * h2d * world * Ke / dt / 1e9,
                       "api": "lc_set_state + lc_step(1) + lc_get_state, all eight component tables, pinned host memory; the three calls "
                              "are dependent (set -> step -> get): nothing can overlap"}
            # The same bytes per frame with the copies overlapped: teacher forcing from a recorded trajectory — the state frame
            # t + 1 starts from is in host memory before frame t ends — through lc_exchange_state, which uploads it on a copy
            # stream of its own beside the frame still executing and beside the download of frame t's result.
            pinned_o = {k: torch.from_numpy(a.copy()).pin_memory() for k, a in st.items()}
            outs = {k: t.numpy() for k, t in pinned_o.items()}
            sim.set_state(**pinned)
            for _ in range(3):
                sim.step(1, sync=False)
                sim.exchange_state(outs, **pinned)
            barrier()
            t0 = time.perf_counter()
            for _ in range(Ke):
                sim.step(1, sync=False)
                sim.exchange_state(outs, **pinned)
            barrier()
            dt = time.perf_counter() - t0
            if dist is not None:
                t = torch.tensor([dt], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = float(t.item())
            e2e = {"value": n_total * Ke / dt, "unit": "agent-steps/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": h2d * world,
                   "steps": Ke, "ms_per_step": 1e3 * dt / Ke, "host_gb_per_s": 2 * h2d * world * Ke / dt / 1e9,
                   "api": "lc_step(1) + lc_exchange_state per frame: all eight component tables down (the frame's result) and all eight "
                          "up (the state the next frame is forced to start from, known beforehand: teacher forcing), pinned host memory; "
                          "the upload rides a copy stream beside the frame and the download",
                   "dependent": e2e_dep}
            del pinned_t, pinned_o

    counters = sim.counters()
    barrier()   # a decomposed crowd's mailboxes are mapped into the peers: nobody frees before everybody is done
    sim.close()

    # the metric's second half: PSO scenario evals/s from the pso4096 batch, sharded over the same N GPUs
    secondary = None
    if not args.no_secondary:
        m = batch_measure("pso4096", 3, 3, rank, world, local, dist, with_e2e=False)
        secondary = {"workload": "pso4096", "scenario_evals_per_sec": m["scenario_evals_per_sec"], "agent_steps_per_sec": m["agent_steps_per_sec"],
                     "ms_per_batch": m["ms_per_step"], "steps": m["steps"], "scaling": "strong", "roofline": m["roofline"],
                     "gpu_launches": m["gpu_launches"], "desc": m["desc"]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        T = host_cores()
        v, ns, dt = crowd_cpu_rate(name, frames=20 if name != "circle1k" else 200, warm=1, threads=T)
        cpu = {"value": v, "unit": "agent-steps/s", "cores": T, "kind": "port", "per_core": v / T, "sample": sample_text(name, T) + f", {dt:.1f} s"}
        r = oref_rate(0, T)
        if r is not None:  # context: the reference's own binary on a workload it can hold
            cpu["oref_scene0"] = {"value": r[0], "unit": "agent-steps/s", "cores": T, "kind": "reference", "scenario_evals_per_sec": r[1],
                                  "sample": f"oracle/_ref/oref, scene 0 (60 agents) to completion, one process per core, {r[3]:.1f} s"}

    if rank == 0:
        line = {"metric": "agent_steps_per_sec", "value": value, "unit": "agent-steps/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak" if (world > 1 and not domain) else "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": name, "desc": WORKLOADS[name]["desc"], "agents": n_total, "frames_per_step": 1,
                           "parallelism": "single crowd, 1 GPU" if world == 1 else
                           (f"one crowd in {world} x strips, one per GPU; halo / migration / candidate exchange through NVLink peer memory "
                            f"inside the frame kernels (rank 0: {n_owned} own + {n_local - n_owned} ghost agents)" if domain
                            else f"{world} independent crowds, one per GPU, no exchange"),
                           "l2": "per-frame working set (2 x 88 B x agents + grid) exceeds the 126 MB L2 at 1M agents; no flush"},
                "counters": counters, "clocks": clk, "e2e": e2e, "e2e_read": e2e_read, "gpu_launches": int(launches), "roofline": roofline,
                "state_hash": state_hash, "timeline": timeline, "secondary": secondary, "cpu_baseline": cpu}
        print(json.dumps(line), file=RESULT_OUT, flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
