#!/usr/bin/env python
"""bench.py — agent-steps/sec of the learn-comms per-frame loop on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload plaza1m|corridor100k|circle1k] [--impl reference]

One "step" = one frame (main.cpp:98-150: demo::run, comm::run, ai::update_agents, physics::simulate, all_sync) of the
whole crowd.  Default workload: BASELINE.json configs[4], the 1M-agent open plaza the north_star's target is quoted
on (it fits one GPU).  `value` = agent-steps/s with the state resident in HBM; `e2e` = the same frame driven through
the C ABI with HOST component tables (lc_set_state from pinned memory -> lc_step(1) -> lc_get_state) inside the
timed region.  `roofline` is for the dominant pass (k_pairs + k_force: pair filter, forces, integration), timed live with CUDA events on the launch stream.
`cpu_baseline` / `--impl reference` time the CPU restatement of the reference loop (oracle/, O-canon with the
uniform grid; the reference's own store cannot hold more than 65 535 entities, Entity.h:12) on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly ONE JSON line.  Libraries write there too (NCCL prints its "NCCL version ..." banner on fd 1),
# so fd 1 is pointed at stderr for the whole run and the result line goes to a private duplicate of the real stdout.
sys.stdout.flush()
RESULT_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)

ALG_BYTES_PER_AGENT_STEP = 112  # 60 B compulsory read + 52 B compulsory write of live agent state
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
    "circle1k": dict(kind="circle", n=1000, radius=80.0, seed=1, wscale=0.1,
                     desc="circle antipodal swap, 1,000 agents, r = 80 m, seed 1", sample=dict(n=1000, radius=80.0)),
}


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


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

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
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
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


def cpu_sample_rate(name, frames, warm=1):
    """Time O-canon (grid search, single thread) on the bounded sample of `name`; returns (agent-steps/s, n, frames)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    pos, goal, md, rng, m_cv, m_f = make_workload(name, sample=True)
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, grid=True)
    o.set_state(**walking_state(pos, goal))
    o.step(warm, fixed=True)
    t0 = time.perf_counter()
    o.step(frames, fixed=True)
    dt = time.perf_counter() - t0
    return pos.shape[0] * frames / dt, pos.shape[0], dt


def sample_text(name):
    w = WORKLOADS[name]
    return f"{w['kind']} sample {w['sample']} of the workload at the same density, O-canon grid search, 1 thread"


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port; the reference's own entity store is limited to
    65 535 entities and its third-party headers are not vendored) on the host cores, one frame of the bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    name = args.workload
    pos, goal, md, rng, m_cv, m_f = make_workload(name, sample=True)
    n = pos.shape[0]
    o = _oracle.CanonSim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, grid=True)
    o.set_state(**walking_state(pos, goal))
    o.step(args.warmup, fixed=True)
    t0 = time.perf_counter()
    o.step(args.steps, fixed=True)
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    line = {"impl": "reference", "metric": "agent_steps_per_sec", "value": v, "unit": "agent-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "desc": WORKLOADS[name]["desc"], "agents": n, "frames_per_step": 1},
            "cpu_baseline": {"value": v, "unit": "agent-steps/s", "cores": 1, "kind": "port", "sample": sample_text(name)},
            "e2e": {"value": v, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), file=RESULT_OUT, flush=True)


def pso_cpu_rate(n_scenarios, n_agents, seed, stride):
    """O-canon brute force (the reference's all-pairs loop), one thread, on every `stride`-th scenario of the batch."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    from learn_comms_b200 import workloads as wl
    pos, goal, m_cv, m_f, md, rng = wl.pso_batch(n_scenarios, n_agents, seed=seed)
    steps, evals = 0, 0
    t0 = time.perf_counter()
    for s in range(0, n_scenarios, stride):
        o = _oracle.CanonSim(pos[s], goal[s], m_cv[s], m_f[s], max_duration=float(md[s]), rng_state=int(rng[s]))
        o.step(10 ** 6)
        steps += o.frame * n_agents
        evals += 1
    dt = time.perf_counter() - t0
    return steps / dt, evals / dt, evals, dt


def run_batch_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    S, n = w["n_scenarios"], w["n_agents"]
    per_step = 4   # scenarios per step: a bounded sample of the batch (one of every S/per_step), ~1 s of CPU each
    tot_steps, tot_evals, tot_dt = 0.0, 0, 0.0
    for k in range(args.warmup + args.steps):
        v, ev, evals, dt = pso_cpu_rate(S, n, w["seed"], S // per_step)
        if k >= args.warmup:
            tot_steps += v * dt; tot_evals += evals; tot_dt += dt
    v = tot_steps / tot_dt
    sample = f"{per_step} of the {S} scenarios per step (every {S // per_step}-th), O-canon brute force, 1 thread"
    line = {"impl": "reference", "metric": "agent_steps_per_sec", "value": v, "unit": "agent-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_dt / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "desc": w["desc"], "scenarios": S, "agents_per_scenario": n},
            "scenario_evals_per_sec": tot_evals / tot_dt,
            "cpu_baseline": {"value": v, "unit": "agent-steps/s", "cores": 1, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), file=RESULT_OUT, flush=True)


def run_batch(args):
    """PSO fitness batch (BASELINE.json configs[2]): one step = every scenario of the batch rolled out to completion in
    ONE kernel launch per GPU.  N > 1: contiguous blocks of scenarios per rank, no data-path collective, scores gathered
    (12 B per scenario); total work is fixed => strong scaling."""
    import torch
    import learn_comms_b200 as lcb
    from learn_comms_b200 import workloads as wl
    from learn_comms_b200 import pso
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    rank, world, local, dist = dist_setup(args.gpus)
    W, K = max(args.warmup, 3), args.steps
    w = WORKLOADS[args.workload]
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
    clocks = ClockSampler(local)
    clocks.start()
    barrier()
    dev_ms, frames_sum, l0 = 0.0, 0, b.launches
    for _ in range(K):
        b.reset()
        out = b.run()
        dev_ms += b.device_ms()   # CUDA events around the rollout kernel on its own stream
        frames_sum += int(out["frames"].sum())
    barrier()
    launches = b.launches - l0
    clk = clocks.stop()
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

    # e2e: host parameter matrices in, host scores out, through the evaluator pso_opt.py would call (reset = H2D of the
    # initial tables + weights, run = rollouts + D2H of the score block), wall clock, max over ranks
    Ke = max(3, min(K, 10))
    ev = pso.BatchEvaluator(pos, goal, md, rng, device=local, rank=rank, world=world, dist=dist)
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

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    achieved = ALG_BYTES_PER_AGENT_STEP * (steps_all / world) / (dev_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(args.workload, {}).get("k_batch_rollout_dram_bytes_per_launch")
    roofline = {"bound": "hbm", "kernel": "k_batch_rollout", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "kernel_share": 1.0,
                "pair_tests_per_sec": pair_tests * K / (dev_ms * 1e-3),
                "note": "state lives in shared memory / registers for the whole rollout: HBM is touched once per scenario, the kernel "
                        "is FP32-issue bound on the all-pairs TTC loop (DESIGN.md); algorithmic bytes = 112 B x agent-steps"}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, evs, k, dtc = pso_cpu_rate(S, n, w["seed"], S // 16)
        cpu = {"value": v, "unit": "agent-steps/s", "cores": 1, "kind": "port", "scenario_evals_per_sec": evs,
               "sample": f"16 of the {S} scenarios (every {S // 16}-th), O-canon brute force, 1 thread, {dtc:.1f} s"}
    if rank == 0:
        line = {"metric": "agent_steps_per_sec", "value": value, "unit": "agent-steps/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "scenario_evals_per_sec": evals,
                "config": {"workload": args.workload, "desc": w["desc"], "scenarios": S, "agents_per_scenario": n,
                           "parallelism": f"{world} GPU(s), contiguous scenario blocks, score gather only",
                           "l2": "no HBM working set: state is staged once into shared memory; nothing to flush"},
                "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu}
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
    ap.add_argument("--replicas", action="store_true", help="N > 1: N independent crowds instead of one decomposed crowd")
    args = ap.parse_args()
    if WORKLOADS[args.workload]["kind"] == "pso":
        return run_batch_reference(args) if args.impl == "reference" else run_batch(args)
    if args.impl == "reference":
        return run_reference(args)

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

    sim.step(W)
    barrier()
    clocks = ClockSampler(local)
    clocks.start()
    l0 = sim.launches
    barrier()
    sim.timer_start()
    sim.step(K, sync=False)
    ms = sim.timer_stop()
    barrier()
    launches = sim.launches - l0
    clk = clocks.stop()
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    n_total = n if domain else world * n   # agents advanced per step by the whole job
    value = n_total * K / (ms * 1e-3)
    n_local = sim.counts()[0] if domain else n
    n_owned = int(sim.owned().sum()) if domain else n

    # roofline leg: per-kernel device time, CUDA events between the launches on the simulation's stream
    prof = sim.step_profiled(20)
    t_force = (prof["k_pairs"] + prof["k_force"]) * 1e-3  # S1-S3a: pair filter + force/integrate, one logical pass over the state
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    achieved = ALG_BYTES_PER_AGENT_STEP * n_owned / t_force / 1e9  # this rank's own agents per k_pairs + k_force launch
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(name, {}).get("k_pairs_k_force_dram_bytes_per_launch")
    frame_ms = sum(prof.values())
    roofline = {"bound": "hbm", "kernel": "k_pairs+k_force", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "kernel_ms": prof, "kernel_share": (prof["k_pairs"] + prof["k_force"]) / frame_ms,
                "note": "pair loop is FP32/issue bound, not HBM bound (DESIGN.md); algorithmic bytes = 112 B x agents per launch"}

    # e2e leg: host-resident component tables, pinned; set_state (H2D) -> one frame -> get_state (D2H) every step
    e2e = None
    if not args.no_e2e and domain:
        # decomposed crowd: the state never leaves the GPUs between frames; the host-facing step is one frame plus the
        # D2H read of the tables a host engine consumes every frame (pos, vel of the rank's own agents)
        cap = sim.counts()[1]
        pinned = dict(ids=torch.zeros(cap, dtype=torch.int32).pin_memory(), pos=torch.zeros((cap, 2)).pin_memory(),
                      vel=torch.zeros((cap, 2)).pin_memory())
        views = {k: v.numpy() for k, v in pinned.items()}
        Ke = max(3, min(K, 50))
        for _ in range(3):
            sim.step(1, sync=False)
            _, k_own = sim.shard(out=views)
        barrier()
        t0 = time.perf_counter()
        for _ in range(Ke):
            sim.step(1, sync=False)
            _, k_own = sim.shard(out=views)
        barrier()
        dt = time.perf_counter() - t0
        d2h = torch.tensor([float(k_own * 20 + 4)], device="cuda")
        dist.all_reduce(d2h)
        t = torch.tensor([dt], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        d2h = int(d2h.item()) // world
        e2e = {"value": n_total * Ke / dt, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": d2h * world,
               "steps": Ke, "ms_per_step": 1e3 * dt / Ke,
               "api": "lc_step(1) + lc_dd_get_shard (ids, pos, vel of the rank's own agents) per rank, pinned host tables; no per-frame H2D: the decomposed state is device-resident"}
    elif not args.no_e2e:
        st = sim.state()
        pinned = {}
        for k, a in st.items():
            t = torch.from_numpy(a.copy()).pin_memory()
            pinned[k] = t.numpy()
        pinned_keep = pinned  # keep tensors alive through the views
        h2d = sum(a.nbytes for a in pinned.values())
        Ke = max(3, min(K, 50))
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
            t = torch.tensor([dt], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": n_total * Ke / dt, "unit": "agent-steps/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": h2d * world,
               "steps": Ke, "ms_per_step": 1e3 * dt / Ke, "api": "lc_set_state + lc_step(1) + lc_get_state, pinned host tables"}
        del pinned_keep

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, ns, dt = cpu_sample_rate(name, frames=30 if name != "circle1k" else 300)
        cpu = {"value": v, "unit": "agent-steps/s", "cores": 1, "kind": "port", "sample": sample_text(name) + f", {dt:.1f} s"}

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
                "counters": sim.counters(), "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu}
        print(json.dumps(line), file=RESULT_OUT, flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
