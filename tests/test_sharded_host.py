"""CPU tests of the host side of the node-sharded mode (zebra_b200/sharded.py): ownership rule,
exchange accounting, and — with two gloo processes — the IPC-handle exchange every rank performs
before connecting.  No CUDA here; the kernels are covered by tests/test_gpu_sharded*.py."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from zebra_b200 import sharded, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ownership_rule_partitions_nodes():
    ids = np.arange(0, 1000)
    for world in (1, 2, 4, 8):
        own = sharded.owner_of(ids, world)
        row = sharded.local_row(ids, world)
        assert own.min() >= 0 and own.max() == world - 1 or world == 1
        # (owner, local row) is a bijection onto the ids, and rows are dense per rank
        assert np.array_equal(row * world + own, ids)
        for r in range(world):
            assert np.array_equal(np.sort(row[own == r]), np.arange(np.count_nonzero(own == r)))
    with pytest.raises(ValueError):
        sharded.owner_of(ids, 3)


def test_exchange_stats_counts_cross_shard_edges():
    st = synth.make_stream("superuser", n_edges=20000, seed=3)
    neg = synth.make_negatives(st, 200)
    for world in (2, 4, 8):
        x = sharded.exchange_stats(st.src, st.dst, neg, world, 20, 2)
        brute = sum(1 for a, b in zip(st.src.tolist(), st.dst.tolist()) if a != b and (a & (world - 1)) != (b & (world - 1)))
        assert x["cross_edges"] == brute
        assert abs(x["cross_fraction"] - (world - 1) / world) < 0.05
        assert x["mailbox_bytes"] == brute * 2 * 2 * 488
        assert x["row_bytes"] == 20000 * 3 * 2 * 320 * (world - 1)
        assert sum(x["updates_per_rank"]) == 2 * 20000 - int(np.count_nonzero(st.src == st.dst))


def test_communities_spread_over_all_ranks():
    st, neg = synth.make_communities("superuser", 4, 200, n_edges=5000)
    assert st.n_edges == 20000 and (np.diff(st.ts) >= 0).all()
    assert st.shape.table_nodes == 4 * synth.SHAPES["superuser"].n_nodes + 1
    assert st.src.max() < st.shape.table_nodes and neg.max() < st.shape.table_nodes and st.src.min() >= 1
    comm = (st.src - 1) // synth.SHAPES["superuser"].n_nodes
    assert np.array_equal((st.dst - 1) // synth.SHAPES["superuser"].n_nodes, comm)      # edges stay inside a community
    for c in range(4):
        own = np.bincount(sharded.owner_of(st.src[comm == c], 4), minlength=4)
        assert own.min() > 0.15 * own.sum()                                              # ... but every community lives on all ranks


WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
from zebra_b200 import sharded
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
mine = bytes([(r * 17 + i) % 256 for i in range(64)])
hs = sharded.gather_handles(mine)
assert len(hs) == w and hs[r] == mine
for q in range(w):
    assert hs[q] == bytes([(q * 17 + i) % 256 for i in range(64)]), q
try:
    sharded.gather_handles(b"short")
    raise SystemExit("short handle accepted")
except ValueError:
    pass
dist.barrier()
print("HANDLES_OK", r, flush=True)
dist.destroy_process_group()
"""


def test_handle_exchange_gloo_world2(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script), ROOT]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("HANDLES_OK") == 2
