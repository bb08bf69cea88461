"""GPU parity of the pipelined epoch path (zebra_tppr_epoch_*, tppr_finder.stream_epoch): every batch
it yields must be byte-identical to what the reference's per-batch streaming_topk returned
(golden recordings) and to the C oracle, whatever the window size, including ragged last batches."""
import numpy as np
import pytest

import oracle
from tests import helpers as H
from zebra_b200 import synth

pytestmark = pytest.mark.gpu


def _finder(*a, **kw):
    from zebra_b200.finder import tppr_finder
    return tppr_finder(*a, **kw)


@pytest.mark.parametrize("name,window", [("stream_wikipedia6k", 4), ("stream_unipartite5k", 128), ("stream_adv_k20", 1),
                                         ("stream_ties", 3), ("stream_nofake", 2), ("stream_k31", 5)])
def test_epoch_matches_reference_recordings(name, window):
    g = H.load(name)
    no_fake = bool(g["no_fake"])
    f = _finder(int(g["num_nodes"]), int(g["k"]), len(g["alphas"]), list(g["alphas"]), list(g["betas"]))
    digs = []
    it = f.stream_epoch(g["src"], g["dst"], None if no_fake else g["neg"], g["ts"], g["eidx"], int(g["bs"]),
                        rows_per_edge=2 if no_fake else 3, window=window)
    for out in it:
        digs.append(H.digest([a for lst in out for a in lst]))
    digs = np.stack(digs)
    assert digs.shape == g["digests"].shape
    bad = np.where((digs != g["digests"]).any(1))[0]
    assert len(bad) == 0, f"first differing batch {bad[:5]} of {len(digs)}"
    assert f.counters()["timeouts"] == 0


@pytest.mark.parametrize("window,bs,n_edges", [(7, 200, 5130), (128, 200, 30000), (2, 333, 4000), (64, 50, 3201)])
def test_epoch_matches_oracle_and_per_batch_calls(window, bs, n_edges):
    st = synth.make_stream("superuser", n_edges=n_edges, seed=31)
    neg = synth.make_negatives(st, bs)
    N, k, al, be = st.shape.table_nodes, 20, [0.1, 0.1], [0.5, 0.95]
    f = _finder(N, k, 2, al, be)
    cpu = oracle.OracleFinder(N, k, 2, al, be)
    nb = 0
    for b, out in enumerate(f.stream_epoch(st.src, st.dst, neg, st.ts, st.eidx, bs, window=window)):
        lo, hi = b * bs, min(n_edges, (b + 1) * bs)
        nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        ref = cpu.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
        for la, lb in zip(out, ref):
            for x, y in zip(la, lb):
                assert x.shape == y.shape and x.dtype == y.dtype
                assert np.array_equal(x, y), f"batch {b}"
        nb += 1
    assert nb == (n_edges + bs - 1) // bs
    # the tables the epoch left behind continue exactly like per-batch calls would
    ids = np.unique(np.concatenate([st.src, st.dst]))[:500]
    for j in range(2):
        for a, (e, n, t, s, nm) in zip(ids, f.state(j, ids)):
            oe, on, ot, os_, onm = cpu.state(j, int(a))
            assert np.array_equal(e, oe) and np.array_equal(n, on) and np.array_equal(t, ot)
            assert np.array_equal(s, os_) and nm == onm


def test_epoch_window_mode_and_update_only():
    st = synth.make_stream("wikipedia", n_edges=4100, seed=9)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes
    f = _finder(N, 20, 2, [0.1, 0.1], [0.5, 0.95])
    per_batch = [tuple(np.stack([a.copy() for a in lst]) for lst in out)
                 for out in f.stream_epoch(st.src, st.dst, neg, st.ts, st.eidx, 200, window=8)]
    f.reset_tppr()
    seen = 0
    for first, nodes, eidx, dt, w in f.stream_epoch(st.src, st.dst, neg, st.ts, st.eidx, 200, window=8, windows=True):
        for q in range(nodes[0].shape[0]):
            for j in range(2):
                assert np.array_equal(nodes[j][q], per_batch[first + q][0][j])
                assert np.array_equal(eidx[j][q], per_batch[first + q][1][j])
                assert np.array_equal(dt[j][q], per_batch[first + q][2][j])
                assert np.array_equal(w[j][q], per_batch[first + q][3][j])
            seen += 1
    assert seen == len(per_batch) == 21
    # rows_per_edge = 0: same final state as compute_val_tppr over the span
    f.reset_tppr()
    assert list(f.stream_epoch(st.src, st.dst, None, st.ts, st.eidx, 200, rows_per_edge=0, window=4)) == []
    g = _finder(N, 20, 2, [0.1, 0.1], [0.5, 0.95])
    g.compute_val_tppr(st.src.astype(np.int64), st.dst.astype(np.int64), st.ts, st.eidx)
    ids = np.unique(st.src)[:300]
    for j in range(2):
        for (e, n, t, s, nm), (e2, n2, t2, s2, nm2) in zip(f.state(j, ids), g.state(j, ids)):
            assert np.array_equal(e, e2) and np.array_equal(n, n2) and np.array_equal(s, s2) and nm == nm2


def test_epoch_error_behaviour():
    """Misuse fails loudly through zebra_last_error, and leaves the handle usable."""
    st = synth.make_stream("wikipedia", n_edges=1000, seed=4)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes
    f = _finder(N, 20, 2, [0.1, 0.1], [0.5, 0.95])
    with pytest.raises(ValueError):
        list(f.stream_epoch(st.src, st.dst[:-1], neg, st.ts, st.eidx, 200))
    with pytest.raises(RuntimeError, match="rows_per_edge"):
        list(f.stream_epoch(st.src, st.dst, neg, st.ts, st.eidx, 200, rows_per_edge=1))
    with pytest.raises(RuntimeError, match="null input"):
        list(f.stream_epoch(st.src, st.dst, None, st.ts, st.eidx, 200, rows_per_edge=3))
    bad = st.src.copy()
    bad[700] = N + 5                                            # out-of-range node in the fourth batch
    it = f.stream_epoch(bad, st.dst, neg, st.ts, st.eidx, 200, window=1)
    with pytest.raises(RuntimeError, match="out of range"):
        list(it)
    # an epoch abandoned half-way releases the handle when the generator is closed
    it = f.stream_epoch(st.src, st.dst, neg, st.ts, st.eidx, 200, window=1)
    next(it)
    with pytest.raises(RuntimeError, match="in progress|part-way"):
        f.streaming_topk(np.concatenate([st.src[:200], st.dst[:200], neg[:200]]).astype(np.int32),
                         np.concatenate([st.ts[:200]] * 3), st.eidx[:200])
    it.close()
    f.reset_tppr()
    cpu = oracle.OracleFinder(N, 20, 2, [0.1, 0.1], [0.5, 0.95])
    for b, out in enumerate(f.stream_epoch(st.src, st.dst, neg, st.ts, st.eidx, 200, window=2)):
        lo, hi = b * 200, (b + 1) * 200
        ref = cpu.streaming_topk(np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32),
                                 np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
        for la, lb in zip(out, ref):
            for x, y in zip(la, lb):
                assert np.array_equal(x, y)
