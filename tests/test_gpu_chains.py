"""Hub chains (tppr_device.cuh "hub chains"): resident chain warps must reproduce the sequential
semantics of utils/util.py:495-574 exactly.  Adversarial streams — few nodes, so most edges join two
chain nodes; self-loops on chain nodes; negatives that are chain nodes or equal an endpoint; replayed
edge ids; duplicate timestamps — through every way of streaming a span, against the C oracle."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

AL, BE = [0.1, 0.1], [0.5, 0.95]


def _finder(*a, **kw):
    from zebra_b200.finder import tppr_finder
    return tppr_finder(*a, **kw)


def _stream(seed, n_nodes, n_edges, hot, p_hot=0.7, p_loop=0.02):
    rng = np.random.default_rng(seed)
    hubs = rng.choice(np.arange(1, n_nodes), size=hot, replace=False)
    pick = lambda: np.where(rng.random(n_edges) < p_hot, hubs[rng.integers(0, hot, n_edges)], rng.integers(1, n_nodes, n_edges))  # noqa: E731
    src, dst = pick(), pick()
    loop = rng.random(n_edges) < p_loop
    dst = np.where(loop, src, dst)
    neg = np.where(rng.random(n_edges) < 0.3, hubs[rng.integers(0, hot, n_edges)], rng.integers(1, n_nodes, n_edges))
    neg = np.where(rng.random(n_edges) < 0.05, src, neg)
    ts = np.floor(np.cumsum(rng.exponential(3.0, n_edges)))
    eidx = np.arange(1, n_edges + 1, dtype=np.int64)
    eidx[rng.random(n_edges) < 0.01] = 1                      # replayed edge ids
    return src.astype(np.int32), dst.astype(np.int32), neg.astype(np.int32), ts, eidx


def _oracle_rows(N, k, src, dst, neg, ts, eidx, bs):
    o = oracle.OracleFinder(N, k, 2, AL, BE)
    rows = []
    for lo in range(0, len(src), bs):
        hi = min(len(src), lo + bs)
        nodes = np.concatenate([src[lo:hi], dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        rows.append(o.streaming_topk(nodes, np.concatenate([ts[lo:hi]] * 3), eidx[lo:hi]))
    return o, rows


def _same_state(f, o, nodes):
    for j in range(2):
        for v, (e, n, t, s, nm) in zip(nodes, f.state(j, nodes)):
            oe, on, ot, os_, onm = o.state(j, v)
            assert np.array_equal(e, oe) and np.array_equal(n, on) and np.array_equal(t, ot), (j, v)
            assert np.array_equal(s, os_) and nm == onm, (j, v)


@pytest.mark.parametrize("seed,n_nodes,hot,k", [(1, 40, 6, 20), (2, 400, 30, 20), (3, 25, 3, 5), (4, 3000, 2, 31)])
def test_chain_streams_match_oracle(seed, n_nodes, hot, k):
    E, bs = 24_000, 200
    src, dst, neg, ts, eidx = _stream(seed, n_nodes, E, hot)
    N = n_nodes + 1
    o, ref = _oracle_rows(N, k, src, dst, neg, ts, eidx, bs)
    nodes = np.arange(1, N)
    # (a) resident plan, several window sizes (a chain spans many launches)
    for per_launch in (120, 16, 3):
        f = _finder(N, k, 2, AL, BE)
        plan = f.make_plan(src, dst, neg, ts, eidx, bs)
        out = plan.run(0, plan.n_batches, per_launch, out_slots=plan.n_batches)
        f.sync_check()
        c = f.counters()
        assert c["chain_steps"] > 0 and c["timeouts"] == 0, c
        for b in range(plan.n_batches):
            got = plan.views(out, b, 0, plan.n_batches)
            for q in range(4):
                for j in range(2):
                    assert np.array_equal(got[q][j].cpu().numpy(), ref[b][q][j]), ("plan", per_launch, b, q, j)
        _same_state(f, o, nodes)
    # (b) pipelined epoch
    f = _finder(N, k, 2, AL, BE)
    for b, got in enumerate(f.stream_epoch(src, dst, neg, ts, eidx, bs, window=32)):
        for q in range(4):
            for j in range(2):
                assert np.array_equal(got[q][j], ref[b][q][j]), ("epoch", b, q, j)
    assert f.counters()["chain_steps"] > 0
    _same_state(f, o, nodes)
    # (c) one long streaming call (rows of the whole span) and the update-only pass
    f = _finder(N, k, 2, AL, BE)
    o2 = oracle.OracleFinder(N, k, 2, AL, BE)
    big = np.concatenate([src, dst, neg]).astype(np.int32)
    a = f.streaming_topk(big, np.concatenate([ts] * 3), eidx)
    b_ = o2.streaming_topk(big, np.concatenate([ts] * 3), eidx)
    assert f.counters()["chain_steps"] > 0
    for q in range(4):
        for j in range(2):
            assert np.array_equal(a[q][j], b_[q][j]), ("stream", q, j)
    f = _finder(N, k, 2, AL, BE)
    f.compute_val_tppr(src.astype(np.int64), dst.astype(np.int64), ts, eidx)
    assert f.counters()["chain_steps"] > 0
    _same_state(f, o, nodes)


def test_chains_off_equals_chains_on(monkeypatch):
    """ZEBRA_CHAINS=0 (every edge through ticket warps only) and the default give identical bytes."""
    src, dst, neg, ts, eidx = _stream(9, 300, 12_000, 8)
    N = 301
    res = []
    for env in ("0", None):
        if env is None:
            monkeypatch.delenv("ZEBRA_CHAINS", raising=False)
        else:
            monkeypatch.setenv("ZEBRA_CHAINS", env)
        f = _finder(N, 20, 2, AL, BE)
        plan = f.make_plan(src, dst, neg, ts, eidx, 200)
        out = plan.run(0, plan.n_batches, 60, out_slots=plan.n_batches)
        f.sync_check()
        res.append((out.cpu().numpy().copy(), f.counters()["chain_steps"]))
    assert res[0][1] == 0 and res[1][1] > 0
    assert np.array_equal(res[0][0], res[1][0])
