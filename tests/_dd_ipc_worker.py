"""Worker of tests/test_gpu_domain_ipc.py: one process per GPU (torchrun), the ranks' mailboxes wired through CUDA IPC
handles (lc_dd_ipc_handle / lc_dd_connect_ipc, exchanged with torch.distributed) — the path bench.py --gpus N uses.
Rank 0 also runs the single-GPU engine and compares the merged state bit for bit.  TEST INFRASTRUCTURE ONLY."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import learn_comms_b200 as lcb  # noqa: E402
from learn_comms_b200 import workloads as wl  # noqa: E402

KEYS = ("pos", "vel", "vfh", "c", "facing", "local_goal", "goal_dist", "num_done")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    rs = np.random.RandomState(11)
    n, w, h = 6000, 160.0, 40.0
    pos = np.stack([rs.uniform(-w / 2, w / 2, n), rs.uniform(-h / 2, h / 2, n)], 1).astype(np.float32)
    goal = np.stack([rs.uniform(-w / 2, w / 2, n), rs.uniform(-h / 2, h / 2, n)], 1).astype(np.float32)
    m_cv, m_f = wl.weights(7, 0.5)
    order = np.random.RandomState(3).permutation(n).astype(np.int32)   # a non-trivial table order on top
    sim = lcb.DomainSim(pos, goal, rank, world, m_cv, m_f, max_duration=12.0, rng_state=4242, device=local, order=order)
    sim.connect_dist(dist)
    ref = lcb.Sim(pos, goal, m_cv, m_f, max_duration=12.0, rng_state=4242, device=local, order=order) if rank == 0 else None
    checked = 0
    for chunk in range(40):
        sim.step(20)
        st, own = sim.state(), sim.owned().astype(bool)
        parts = [None] * world
        dist.all_gather_object(parts, ({k: v[own] for k, v in st.items()}, np.flatnonzero(own)))
        fin = sim.finished
        if rank == 0:
            ref.step(20)
            full = ref.state()
            seen = np.zeros(n, np.int32)
            for sub, ids in parts:
                seen[ids] += 1
                for k in KEYS:
                    a, b = sub[k], full[k][ids]
                    same = np.array_equal(a, b, equal_nan=True) if a.dtype.kind == "f" else np.array_equal(a, b)
                    assert same, f"frame {20 * (chunk + 1)}: {k} differs on the IPC path"
            assert (seen == 1).all(), "ownership is not a partition"
            checked += 1
        if fin:
            break
    sc = sim.scores_dist(dist)
    if rank == 0:
        assert ref.finished and sim.finished and ref.frame == sim.frame, (ref.frame, sim.frame)
        rsc = ref.scores()
        for k in rsc:
            assert sc[k] == rsc[k] or (np.isnan(sc[k]) and np.isnan(rsc[k])), (k, sc, rsc)
        print(f"DD_IPC_OK world={world} checked={checked} frames={sim.frame} devices={torch.cuda.device_count()}", flush=True)
    dist.barrier()
    sim.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
