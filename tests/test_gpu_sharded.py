"""GPU parity of the node-sharded span kernel (zebra_b200/sharded.py, tppr_shard_device.cuh).

The sharded protocol (ownership, mailbox pushes, rows written into every rank's output region) is
exercised here on ONE GPU: `world` handles in one process, each capped to 1/world of the persistent
grid so that all ranks' kernels are co-resident, each on its own stream.  Every rank's output must be
byte-identical to the C oracle's per-batch results (and so to the single-GPU path), and the owned
vectors must equal the oracle's final state bit for bit.  tests/test_gpu_sharded_mp.py runs the same
check with one process per GPU over CUDA IPC when the box has more than one GPU."""
import numpy as np
import pytest

import oracle
from zebra_b200 import synth

pytestmark = pytest.mark.gpu

AL, BE = [0.1, 0.1], [0.5, 0.95]


def _run_sharded(world, st, neg, bs, k, window, rows_per_edge=3, row_mode=0):
    import torch
    from zebra_b200.sharded import ShardedFinder
    N = st.shape.table_nodes
    n_batches = (st.n_edges + bs - 1) // bs
    out_bytes = n_batches * 2 * 16 * rows_per_edge * bs * k
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    fs = [ShardedFinder(N, k, 2, AL, BE, r, world, window * bs, out_bytes) for r in range(world)]
    for f in fs:
        f.set_max_blocks(max(1, (2 * sms) // world))
        f.set_row_mode(row_mode)
    ShardedFinder.connect_local(fs)
    plans = [f.make_plan(st.src, st.dst, neg if rows_per_edge == 3 else None, st.ts, st.eidx, bs, rows_per_edge) for f in fs]
    streams = [torch.cuda.Stream() for _ in fs]
    torch.cuda.synchronize()
    for f, p, s in zip(fs, plans, streams):
        with torch.cuda.stream(s):
            p.run(0, n_batches, window, out_slots=n_batches)
    for f, s in zip(fs, streams):
        with torch.cuda.stream(s):
            f.sync_check()
    torch.cuda.synchronize()
    return fs, plans, n_batches


@pytest.mark.parametrize("world,name,n_edges,bs,window", [(2, "wikipedia", 6000, 200, 8), (4, "superuser", 5000, 200, 4),
                                                         (2, "reddit", 3100, 100, 64), (8, "superuser", 3000, 200, 2)])
def test_sharded_matches_oracle(world, name, n_edges, bs, window):
    st = synth.make_stream(name, n_edges=n_edges, seed=77)
    neg = synth.make_negatives(st, bs)
    N, k = st.shape.table_nodes, 20
    fs, plans, n_batches = _run_sharded(world, st, neg, bs, k, window)
    cpu = oracle.OracleFinder(N, k, 2, AL, BE)
    for b in range(n_batches):
        lo, hi = b * bs, min(n_edges, (b + 1) * bs)
        nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        ref = cpu.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
        for f, p in zip(fs, plans):
            got = p.views(f.out, b, 0, n_batches)
            for la, lb in zip(got, ref):
                for x, y in zip(la, lb):
                    assert np.array_equal(x.cpu().numpy(), y), f"rank {f.rank} batch {b}"
    ids = np.unique(np.concatenate([st.src, st.dst]))
    for f in fs:
        mine = ids[(ids & (world - 1)) == f.rank][:400]
        for j in range(2):
            for a, (e, n, t, s, nm) in zip(mine, f.state(j, mine)):
                oe, on, ot, os_, onm = cpu.state(j, int(a))
                assert np.array_equal(e, oe) and np.array_equal(n, on) and np.array_equal(t, ot)
                assert np.array_equal(s, os_), "fp64 scores must be bit-identical"
                assert nm == onm
        assert f.counters()["timeouts"] == 0
    with pytest.raises(RuntimeError):
        fs[0].state(0, np.array([1]))      # node 1 lives on rank 1


def test_sharded_adversarial_self_loops_and_ties():
    """Self-loops, negatives equal to an endpoint, exact score ties (alpha=beta=0.5), repeated pairs."""
    rng = np.random.default_rng(3)
    E, N, bs, k = 2400, 64, 100, 8
    src = rng.integers(1, N, E).astype(np.int32)
    dst = rng.integers(1, N, E).astype(np.int32)
    dst[::17] = src[::17]
    neg = rng.integers(1, N, E).astype(np.int32)
    neg[::11] = src[::11]
    neg[5::13] = dst[5::13]
    ts = np.floor(np.cumsum(rng.exponential(3.0, E)))
    eidx = np.arange(1, E + 1, dtype=np.int64)
    shape = synth.StreamShape("adv", N - 1, 0, E, 0, 1.0, 0.1, 0.0)
    st = synth.Stream(shape, src, dst, ts, eidx, np.unique(dst))
    global AL, BE
    old = (AL, BE)
    AL, BE = [0.5, 0.1], [0.5, 0.5]
    try:
        fs, plans, n_batches = _run_sharded(2, st, neg, bs, k, 3)
        cpu = oracle.OracleFinder(N, k, 2, AL, BE)
        for b in range(n_batches):
            lo, hi = b * bs, min(E, (b + 1) * bs)
            nodes = np.concatenate([src[lo:hi], dst[lo:hi], neg[lo:hi]]).astype(np.int32)
            ref = cpu.streaming_topk(nodes, np.concatenate([ts[lo:hi]] * 3), eidx[lo:hi])
            for f, p in zip(fs, plans):
                got = p.views(f.out, b, 0, n_batches)
                for la, lb in zip(got, ref):
                    for x, y in zip(la, lb):
                        assert np.array_equal(x.cpu().numpy(), y), f"rank {f.rank} batch {b}"
    finally:
        AL, BE = old


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_row_mode_delivers_each_batch_to_one_rank(world):
    """row_mode 1: the rows of batch b land on rank b & (world-1) only; the union is the full result."""
    import torch
    n_edges, bs = 4200, 200
    st = synth.make_stream("superuser", n_edges=n_edges, seed=78)
    neg = synth.make_negatives(st, bs)
    N, k = st.shape.table_nodes, 20
    fs, plans, n_batches = _run_sharded(world, st, neg, bs, k, 4, row_mode=1)
    cpu = oracle.OracleFinder(N, k, 2, AL, BE)
    for b in range(n_batches):
        lo, hi = b * bs, min(n_edges, (b + 1) * bs)
        nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        ref = cpu.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
        for f, p in zip(fs, plans):
            got = p.views(f.out, b, 0, n_batches)
            if f.rank == (b & (world - 1)):
                for la, lb in zip(got, ref):
                    for x, y in zip(la, lb):
                        assert np.array_equal(x.cpu().numpy(), y), f"rank {f.rank} batch {b}"
            else:                                  # never written: still the zero fill of the arena
                assert all(int(torch.count_nonzero(x)) == 0 for la in got for x in la), f"rank {f.rank} batch {b}"


def test_sharded_error_behaviour():
    from zebra_b200.sharded import ShardedFinder
    st = synth.make_stream("wikipedia", n_edges=1000, seed=4)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes
    with pytest.raises(RuntimeError, match="world must be"):
        ShardedFinder(N, 20, 2, AL, BE, 0, 3, 1000, 1 << 20)
    f = ShardedFinder(N, 20, 2, AL, BE, 0, 2, 400, 1 << 20)
    plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, 200)
    with pytest.raises(RuntimeError, match="shard_connect"):
        plan.run(0, 1, 1)                                       # peers not connected yet
    with pytest.raises(ValueError, match="window"):
        plan.run(0, 4, 4)                                       # 800 edges > the 400-edge mailbox window
    g = ShardedFinder(N, 20, 2, AL, BE, 1, 2, 400, 1 << 20)
    ShardedFinder.connect_local([f, g])
    with pytest.raises(RuntimeError, match="exceeds the shared output region"):
        plan.run(0, 5, 1, out_slots=5, out_offset=(1 << 20) - 4096)
    with pytest.raises(RuntimeError, match="stream through plans"):
        nodes = np.concatenate([st.src[:1], st.dst[:1]]).astype(np.int32)
        f._check(f._L.zebra_tppr_stream(f._h, nodes.ctypes.data, st.ts[:1].ctypes.data, st.eidx[:1].ctypes.data, 1, 2, -1,
                                        None, None, None))


@pytest.mark.parametrize("chains", ["0", "1"])
@pytest.mark.parametrize("world,name,n_edges,bs,window,row_mode", [(2, "reddit", 9000, 200, 4, 1), (4, "superuser", 7000, 100, 8, 1),
                                                                   (2, "wikipedia", 5000, 200, 6, 0)])
def test_sharded_pipelined_epoch_matches_oracle(world, name, n_edges, bs, window, row_mode, chains, monkeypatch):
    """zebra_tppr_epoch_* on node-sharded handles: windows staged by worker threads, rows copied out once every
    rank has finished the window, output halves re-used only after their consumer has drained them."""
    import torch
    monkeypatch.setenv("ZEBRA_SHARD_CHAINS", chains)
    from zebra_b200.sharded import ShardedFinder
    st = synth.make_stream(name, n_edges=n_edges, seed=78)
    neg = synth.make_negatives(st, bs)
    N, k = st.shape.table_nodes, 20
    n_batches = (n_edges + bs - 1) // bs
    slot = 2 * 16 * 3 * bs * k
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    fs = [ShardedFinder(N, k, 2, AL, BE, r, world, window * bs, 2 * window * slot) for r in range(world)]
    for f in fs:
        f.set_max_blocks(max(1, (2 * sms) // world))
        f.set_row_mode(row_mode)
    ShardedFinder.connect_local(fs)
    streams = [torch.cuda.Stream() for _ in fs]
    for rep in range(2):
        eps = []
        for f, s in zip(fs, streams):
            f.reset_tppr()
            with torch.cuda.stream(s):
                eps.append(iter(f.epoch(st.src, st.dst, neg, st.ts, st.eidx, bs, window=window)))
        cpu = oracle.OracleFinder(N, k, 2, AL, BE)
        refs = []
        for b in range(n_batches):
            lo, hi = b * bs, min(n_edges, (b + 1) * bs)
            nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
            refs.append(cpu.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi]))
        # the ranks' iterators are advanced in turn (one process stands in for all ranks)
        seen = [[] for _ in fs]
        live = list(range(world))
        while live:
            for r in list(live):
                item = next(eps[r], None)
                if item is None:
                    live.remove(r)
                    continue
                b, nd, ei, dt, w = item
                for got, want in zip((nd, ei, dt, w), refs[b]):
                    for x, y in zip(got, want):
                        assert np.array_equal(x, y), (rep, r, b)
                seen[r].append(b)
        for r in range(world):
            want = list(range(r, n_batches, world)) if row_mode == 1 else list(range(n_batches))
            assert seen[r] == want, (r, seen[r][:6])
    # the same pass consumed window by window
    for f in fs:
        f.reset_tppr()
    eps = []
    for f, s_ in zip(fs, streams):
        with torch.cuda.stream(s_):
            eps.append(iter(f.epoch(st.src, st.dst, neg, st.ts, st.eidx, bs, window=window, windows=True)))
    live = list(range(world))
    got_batches = [[] for _ in fs]
    while live:
        for r in list(live):
            item = next(eps[r], None)
            if item is None:
                live.remove(r)
                continue
            idxs, nd, ei, dt, w = item
            for q, b in enumerate(idxs):
                for got, want in zip((nd, ei, dt, w), refs[b]):
                    for x, y in zip(got, want):
                        assert np.array_equal(x[q], y), ("windows", r, b)
                got_batches[r].append(b)
    for r in range(world):
        assert got_batches[r] == (list(range(r, n_batches, world)) if row_mode == 1 else list(range(n_batches)))
    for f in fs:
        assert f.counters()["timeouts"] == 0


@pytest.mark.parametrize("world,seed,n_nodes,hot", [(2, 11, 60, 6), (4, 12, 500, 20), (2, 13, 30, 3)])
def test_sharded_chain_warps_match_oracle(world, seed, n_nodes, hot, monkeypatch):
    """Hub chains in the node-sharded kernel: the owner rank of a hot node walks its edges with a resident chain
    warp; partners on other ranks exchange vectors with it through the NVLink mailboxes.  Adversarial stream
    (most edges between hot nodes, self-loops, negatives on hot nodes) through plans and the pipelined epoch."""
    import torch
    from tests.test_gpu_chains import _stream
    from zebra_b200.sharded import ShardedFinder
    monkeypatch.setenv("ZEBRA_SHARD_CHAINS", "1")          # off by default on sharded handles
    E, bs, k, window = 16_000, 200, 20, 16
    src, dst, neg, ts, eidx = _stream(seed, n_nodes, E, hot)
    N = n_nodes + 1
    n_batches = E // bs
    slot = 2 * 16 * 3 * bs * k
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    cpu = oracle.OracleFinder(N, k, 2, AL, BE)
    refs = []
    for b in range(n_batches):
        lo, hi = b * bs, (b + 1) * bs
        nodes = np.concatenate([src[lo:hi], dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        refs.append(cpu.streaming_topk(nodes, np.concatenate([ts[lo:hi]] * 3), eidx[lo:hi]))
    fs = [ShardedFinder(N, k, 2, AL, BE, r, world, window * bs, n_batches * slot) for r in range(world)]
    for f in fs:
        f.set_max_blocks(max(1, (2 * sms) // world))
    ShardedFinder.connect_local(fs)
    streams = [torch.cuda.Stream() for _ in fs]
    plans = [f.make_plan(src, dst, neg, ts, eidx, bs) for f in fs]
    torch.cuda.synchronize()
    for f, p, s_ in zip(fs, plans, streams):
        with torch.cuda.stream(s_):
            p.run(0, n_batches, window, out_slots=n_batches)
    for f, s_ in zip(fs, streams):
        with torch.cuda.stream(s_):
            f.sync_check()
    torch.cuda.synchronize()
    assert sum(f.counters()["chain_steps"] for f in fs) > 0
    for b in range(n_batches):
        for f, p in zip(fs, plans):
            got = p.views(f.out, b, 0, n_batches)
            for la, lb in zip(got, refs[b]):
                for x, y in zip(la, lb):
                    assert np.array_equal(x.cpu().numpy(), y), f"plan: rank {f.rank} batch {b}"
    ids = np.arange(1, N)
    for f in fs:
        mine = ids[(ids & (world - 1)) == f.rank]
        for j in range(2):
            for a, (e, n, t, s, nm) in zip(mine, f.state(j, mine)):
                oe, on, ot, os_, onm = cpu.state(j, int(a))
                assert np.array_equal(e, oe) and np.array_equal(s, os_) and nm == onm, (f.rank, j, a)
    # the pipelined epoch over the same stream (row mode 1)
    for f in fs:
        f.set_row_mode(1)
        f.reset_tppr()
    eps = []
    for f, s_ in zip(fs, streams):
        with torch.cuda.stream(s_):
            eps.append(iter(f.epoch(src, dst, neg, ts, eidx, bs, window=window)))
    live = list(range(world))
    count = [0] * world
    while live:
        for r in list(live):
            item = next(eps[r], None)
            if item is None:
                live.remove(r)
                continue
            b, nd, ei, dt, w = item
            for got, want in zip((nd, ei, dt, w), refs[b]):
                for x, y in zip(got, want):
                    assert np.array_equal(x, y), ("epoch", r, b)
            count[r] += 1
    assert sum(count) == n_batches
    for f in fs:
        assert f.counters()["timeouts"] == 0
