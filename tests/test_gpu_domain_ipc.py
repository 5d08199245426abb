"""The domain decomposition through the path bench.py actually uses: ONE PROCESS PER GPU, mailboxes mapped into the
peers with CUDA IPC handles (lc_dd_ipc_handle / lc_dd_connect_ipc), ranks on different devices.  Needs >= 2 GPUs (the
single-GPU box of the round-end run skips it; profiles/ holds the log of a 2-GPU box)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 4])
def test_domain_ipc_processes_bit_exact(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, the box has {torch.cuda.device_count()}")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "_dd_ipc_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=dict(os.environ, LC_DD_TIMEOUT_MS="60000"))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert f"DD_IPC_OK world={world}" in r.stdout
