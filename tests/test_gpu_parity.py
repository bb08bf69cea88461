"""GPU parity: the CUDA span kernel (through the C ABI and the tppr_finder drop-in) against
(1) recordings of the reference itself (tests/golden, bit-exact incl. exact score ties) and
(2) the C oracle on larger seeded streams, plus size-independent properties at full size."""
import numpy as np
import pytest

import oracle
from tests import helpers as H
from zebra_b200 import synth

pytestmark = pytest.mark.gpu


def _finder(*a, **kw):
    from zebra_b200.finder import tppr_finder
    return tppr_finder(*a, **kw)


def _check_state(f, g):
    for j in range(f.n_tppr):
        st = f.state(j, g["state_nodes"])
        for a, (e, n, t, s, nm) in enumerate(st):
            c = int(g["st_cnt"][j, a])
            assert len(e) == c
            assert np.array_equal(e, g["st_eidx"][j, a, :c]) and np.array_equal(n, g["st_node"][j, a, :c])
            assert np.array_equal(t, g["st_ts"][j, a, :c])
            assert np.array_equal(s, g["st_score"][j, a, :c]), "fp64 scores must be bit-identical"
            assert nm == g["st_norm"][j, a]


@pytest.mark.parametrize("name", H.STREAM_CASES)
def test_golden_stream_bit_exact(name):
    g = H.load(name)
    f = _finder(int(g["num_nodes"]), int(g["k"]), len(g["alphas"]), list(g["alphas"]), list(g["betas"]))
    digs, outs = H.replay(f, g)
    bad = np.where((digs != g["digests"]).any(1))[0]
    assert len(bad) == 0, f"first differing batch {bad[:5]} of {len(digs)}"
    for key in g.files:
        if key.startswith("o_node_"):
            b = int(key.split("_")[-1])
            for q, nm in enumerate(("o_node", "o_eidx", "o_dt", "o_w")):
                assert np.array_equal(outs[b][q], g[f"{nm}_{b}"])
    _check_state(f, g)
    assert f.counters()["timeouts"] == 0


def test_protocol_aliasing_matches_reference():
    g = H.load("protocol_alias")
    src, dst, neg, ts, eidx = g["src"], g["dst"], g["neg"], g["ts"], g["eidx"]
    f = _finder(int(g["num_nodes"]), int(g["k"]), 2, list(g["alphas"]), list(g["betas"]))

    def batch(lo, hi):
        return (np.concatenate([src[lo:hi], dst[lo:hi], neg[lo:hi]]).astype(np.int32),
                np.concatenate([ts[lo:hi]] * 3), eidx[lo:hi].astype(np.int64))

    S = lambda o: np.stack([np.stack(x) for x in o]).astype(np.float64)  # noqa: E731
    f.streaming_topk(*batch(0, 200))
    f.reset_tppr()
    f.compute_val_tppr(src[:1000].astype(np.int64), dst[:1000].astype(np.int64), ts[:1000], eidx[:1000].astype(np.int64))
    bk = f.backup_tppr()
    assert np.array_equal(S(f.streaming_topk(*batch(1000, 1200))), g["val1"])
    bk2 = f.backup_tppr()
    f.restore_tppr(bk)
    assert np.array_equal(S(f.streaming_topk(*batch(1000, 1200))), g["nn_val1"])
    f.restore_tppr(bk2)
    f.reset_tppr()
    f.streaming_topk(*batch(0, 200))
    f.reset_tppr()
    f.restore_val_tppr()
    assert np.array_equal(S(f.streaming_topk(*batch(1000, 1200))), g["val2"])
    s0 = f.single_streaming_topk(*batch(1200, 1400), 1)
    assert np.array_equal(np.stack([np.asarray(x, np.float64) for x in s0]), g["single"])
    _check_state(f, g)


def test_strict_snapshots_roll_back():
    st = synth.make_stream("wikipedia", n_edges=1200, seed=5)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes
    f = _finder(N, 20, 2, [0.1, 0.1], [0.5, 0.95], strict_snapshots=True)
    o = oracle.OracleFinder(N, 20, 2, [0.1, 0.1], [0.5, 0.95], strict_snapshots=True)

    def batch(lo, hi):
        return (np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32),
                np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
    for x in (f, o):
        x.streaming_topk(*batch(0, 400))
    bf, bo = f.backup_tppr(), o.backup_tppr()
    r1 = [f.streaming_topk(*batch(400, 600)), o.streaming_topk(*batch(400, 600))]
    f.restore_tppr(bf); o.restore_tppr(bo)
    r2 = [f.streaming_topk(*batch(400, 600)), o.streaming_topk(*batch(400, 600))]
    for q in range(4):
        for j in range(2):
            assert np.array_equal(r1[0][q][j], r1[1][q][j]) and np.array_equal(r2[0][q][j], r2[1][q][j])
            assert np.array_equal(r1[0][q][j], r2[0][q][j]), "strict restore must roll the table back"


@pytest.mark.parametrize("name,n_edges,bs", [("wikipedia", 40000, 200), ("reddit", 30000, 200), ("superuser", 40000, 200),
                                            ("wikitalk", 60000, 1000), ("reddit", 20000, 7)])
def test_seeded_stream_vs_oracle(name, n_edges, bs):
    st = synth.make_stream(name, n_edges=n_edges)
    neg = synth.make_negatives(st, bs)
    N = st.shape.table_nodes
    al, be = [0.1, 0.1], [0.5, 0.95]
    f = _finder(N, 20, 2, al, be)
    o = oracle.OracleFinder(N, 20, 2, al, be)
    for lo in range(0, n_edges, bs):
        hi = min(n_edges, lo + bs)
        nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        ts = np.concatenate([st.ts[lo:hi]] * 3)
        a = f.streaming_topk(nodes, ts, st.eidx[lo:hi])
        b = o.streaming_topk(nodes, ts, st.eidx[lo:hi])
        for q in range(4):
            for j in range(2):
                assert np.array_equal(a[q][j], b[q][j]), (name, lo, q, j)
    touched = np.unique(np.concatenate([st.src, st.dst]))[:500]
    for j in range(2):
        for v, (e, n, t, s, nm) in zip(touched, f.state(j, touched)):
            oe, on, ot, os_, onm = o.state(j, v)
            assert np.array_equal(e, oe) and np.array_equal(n, on) and np.array_equal(s, os_) and nm == onm
    assert f.counters()["timeouts"] == 0


def test_update_only_long_span_vs_oracle_and_batched():
    """compute_val_tppr over a multi-chunk span == the oracle == the same edges fed in small batches."""
    st = synth.make_stream("superuser", n_edges=600_000)
    N = st.shape.table_nodes
    al, be = [0.1, 0.1], [0.5, 0.95]
    f = _finder(N, 20, 2, al, be)
    o = oracle.OracleFinder(N, 20, 2, al, be)
    f.compute_val_tppr(st.src, st.dst, st.ts, st.eidx)
    o.compute_val_tppr(st.src, st.dst, st.ts, st.eidx)
    rng = np.random.default_rng(0)
    deg = np.bincount(np.concatenate([st.src, st.dst]), minlength=N)
    sample = np.unique(np.concatenate([np.argsort(deg)[-200:], rng.choice(np.nonzero(deg)[0], 800)]))
    for j in range(2):
        for v, (e, n, t, s, nm) in zip(sample, f.state(j, sample)):
            oe, on, ot, os_, onm = o.state(j, v)
            assert np.array_equal(e, oe) and np.array_equal(n, on) and np.array_equal(s, os_) and nm == onm
    g = _finder(N, 20, 2, al, be)
    for lo in range(0, 600_000, 50_000):
        hi = lo + 50_000
        g.compute_val_tppr(st.src[lo:hi], st.dst[lo:hi], st.ts[lo:hi], st.eidx[lo:hi])
    for j in range(2):
        for a, b in zip(f.state(j, sample), g.state(j, sample)):
            assert all(np.array_equal(x, y) for x, y in zip(a[:4], b[:4])) and a[4] == b[4]


def test_full_size_properties_wikitalk():
    """BASELINE.json configs[4] at full size: properties that do not need the oracle."""
    st = synth.make_stream("wikitalk")
    N = st.shape.table_nodes
    be = [0.5, 0.95]
    f = _finder(N, 20, 2, [0.1, 0.1], be)
    f.compute_val_tppr(st.src, st.dst, st.ts, st.eidx)
    deg = np.bincount(np.concatenate([st.src, st.dst]), minlength=N)
    loops = np.bincount(st.src[st.src == st.dst], minlength=N)
    deg = deg - loops                                   # a self-loop updates its node once
    rng = np.random.default_rng(1)
    sample = np.unique(np.concatenate([np.argsort(deg)[-100:], rng.integers(0, N, 3000)]))
    for j in range(2):
        for v, (e, n, t, s, nm) in zip(sample, f.state(j, sample)):
            d = int(deg[v])
            # norm = sum_{z=1..deg} beta^z, evaluated with the reference's recurrence
            ref = 0.0
            for _ in range(min(d, 5000)):
                ref = ref * be[j] + be[j]
            if d <= 5000:
                assert nm == ref
            assert len(e) == min(20, len(e)) and (d == 0) == (len(e) == 0)
            if len(e):
                assert np.all(s > 0) and np.all(e >= 1) and np.all(e <= st.n_edges)
                assert np.all(st.ts[e - 1] == t), "stored timestamps are the edges' own"
                assert len(set(zip(e.tolist(), n.tolist()))) == len(e), "keys are unique"
                assert s.sum() <= 1.0 + 1e-9
            if len(e) == 20 and d > 25:
                assert np.all(np.diff(s) >= 0), "after a prune slots are in ascending score order"
    assert f.counters()["timeouts"] == 0


def test_errors_are_loud():
    f = _finder(50, 5, 1, [0.1], [0.5])
    with pytest.raises(RuntimeError):
        f.streaming_topk(np.array([1, 2, 60], np.int32), np.zeros(3), np.array([1], np.int64))
    with pytest.raises(RuntimeError):
        f.streaming_topk(np.array([1, 2, 3], np.int32), np.zeros(3), np.array([2**40], np.int64))
    # the handle still works after a rejected call
    o = f.streaming_topk(np.array([1, 2, 3], np.int32), np.zeros(3), np.array([1], np.int64))
    assert o[0][0].shape == (3, 5)
    # empty batch
    o = f.streaming_topk(np.zeros(0, np.int32), np.zeros(0), np.zeros(0, np.int64))
    assert o[0][0].shape == (0, 5)


@pytest.mark.parametrize("per_launch", [1, 7, 64])
def test_resident_plan_equals_per_batch_calls(per_launch):
    """A resident plan replayed with several batches per launch == one streaming_topk call per batch."""
    import torch
    st = synth.make_stream("reddit", n_edges=12_345)
    bs = 200
    neg = synth.make_negatives(st, bs)
    N = st.shape.table_nodes
    al, be = [0.1, 0.1], [0.5, 0.95]
    f = _finder(N, 20, 2, al, be)
    o = oracle.OracleFinder(N, 20, 2, al, be)
    plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, bs)
    nb = plan.n_batches
    assert nb == (12_345 + bs - 1) // bs
    out = plan.run(0, nb, per_launch, out_slots=nb)
    f.sync_check()
    for b in range(nb):
        lo, hi = b * bs, min(st.n_edges, (b + 1) * bs)
        nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        ref = o.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
        got = plan.views(out, b, 0, nb)
        for q in range(4):
            for j in range(2):
                assert np.array_equal(got[q][j].cpu().numpy(), ref[q][j]), (b, q, j)
    # the handle keeps working with ordinary calls after the plan finished
    lo, hi = 0, 200
    nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
    a = f.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
    r = o.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
    assert all(np.array_equal(a[q][j], r[q][j]) for q in range(4) for j in range(2))
    assert f.counters()["timeouts"] == 0
