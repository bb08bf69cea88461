"""Node-sharded T-PPR with one process per GPU (CUDA IPC mailboxes over NVLink): spawns torchrun on
all visible GPUs (2, 4 or 8) and lets every rank check its result against the C oracle.  Skipped on a
one-GPU box, where tests/test_gpu_sharded.py covers the same kernel with several ranks on one device."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("name,n_edges,bs,window", [("superuser", 20000, 200, 16), ("reddit", 8000, 200, 64)])
def test_sharded_multi_process(name, n_edges, bs, window):
    import torch
    n = torch.cuda.device_count()
    world = 8 if n >= 8 else 4 if n >= 4 else 2 if n >= 2 else 0
    if not world:
        pytest.skip("needs at least 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "shard_worker.py"), name, str(n_edges), str(bs), str(window)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("SHARD_OK") == world, r.stdout[-3000:]
