"""Announced passes (tppr_finder.schedule_pass / zebra_b200.batches.PassSchedule):
the model's per-batch queries answered from rows computed a window ahead must equal per-call streaming — against
the oracle, for hits, for the roll-back of a mismatching call, and through eval_edge_prediction."""
import types

import numpy as np
import pytest

import oracle
from zebra_b200 import synth

pytestmark = pytest.mark.gpu

AL, BE = [0.1, 0.1], [0.5, 0.95]


def _finder(N, k=20):
    from zebra_b200.finder import tppr_finder
    return tppr_finder(N, k, 2, AL, BE)


def _args(st, neg, bs, b):
    lo, hi = b * bs, min(st.n_edges, (b + 1) * bs)
    return (np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32),
            np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])


def _same(a, b):
    return all(np.array_equal(np.asarray(a[q][j].cpu() if hasattr(a[q][j], "cpu") else a[q][j]), b[q][j])
               for q in range(4) for j in range(2))


def _state_equal(f, o, nodes):
    for j in range(2):
        got = f.state(j, nodes)
        for v, (e, nd, t, s, nm) in zip(nodes, got):
            re, rn, rt, rs, rnm = o.state(j, int(v))
            assert np.array_equal(e, re) and np.array_equal(nd, rn) and np.array_equal(t, rt) and np.array_equal(s, rs)
            assert nm == rnm


@pytest.mark.parametrize("window", [1, 7, 64])
def test_announced_pass_equals_per_call_and_oracle(window):
    st = synth.make_stream("reddit", n_edges=4_321)          # ragged last batch (121 edges)
    bs = 200
    neg = synth.make_negatives(st, bs)
    N = st.shape.table_nodes
    f, o = _finder(N), oracle.OracleFinder(N, 20, 2, AL, BE)
    warm = 5                                                  # some per-call history before the pass
    for b in range(warm):
        a = _args(st, neg, bs, b)
        assert _same(f.streaming_topk(*a), o.streaming_topk(*a))
    lo = warm * bs
    sp = f.schedule_pass(st.src[lo:], st.dst[lo:], neg[lo:], st.ts[lo:], st.eidx[lo:], bs, window)
    nb = sp.n_batches
    for b in range(nb):
        a = _args(st, neg, bs, warm + b)
        ref = o.streaming_topk(*a)
        # alternate the numpy-returning call and the device-resident one (what the fused gather consumes)
        got = f.streaming_topk(*a) if b % 2 else f.streaming_topk_device(*a)
        assert _same(got, ref), b
    assert sp.hits == nb and sp.launches == -(-nb // window) and not sp.rolled_back
    assert f._sched is None
    f.sync_check()
    _state_equal(f, o, np.unique(np.concatenate([st.src[:400], st.dst[-400:]])).astype(np.int64)[:300])
    assert f.counters()["timeouts"] == 0


def test_mismatching_call_rolls_back_and_state_calls_settle():
    st = synth.make_stream("wikipedia", n_edges=6_000)
    bs = 200
    neg = synth.make_negatives(st, bs)
    N = st.shape.table_nodes
    f, o = _finder(N), oracle.OracleFinder(N, 20, 2, AL, BE)
    probe = np.unique(np.concatenate([st.src, st.dst])).astype(np.int64)[:400]
    sp = f.schedule_pass(st.src, st.dst, neg, st.ts, st.eidx, bs, 8)
    for b in range(11):                                       # window 2 (batches 8..15) is in flight, 3 of it consumed
        a = _args(st, neg, bs, b)
        assert _same(f.streaming_topk(*a), o.streaming_topk(*a))
    # the model brings something else: batch 11 with other negatives
    nodes, ts, eidx = _args(st, neg, bs, 11)
    nodes = nodes.copy()
    nodes[2 * bs:] = nodes[2 * bs:][::-1]
    assert _same(f.streaming_topk(nodes, ts, eidx), o.streaming_topk(nodes, ts, eidx))
    assert sp.rolled_back and sp.hits == 11 and f._sched is None
    for b in range(12, 15):                                   # and carries on per call
        a = _args(st, neg, bs, b)
        assert _same(f.streaming_topk(*a), o.streaming_topk(*a))
    _state_equal(f, o, probe)
    # a state call in the middle of an announced pass settles it first
    lo = 15 * bs
    sp2 = f.schedule_pass(st.src[lo:], st.dst[lo:], neg[lo:], st.ts[lo:], st.eidx[lo:], bs, 8)
    for b in range(15, 18):
        a = _args(st, neg, bs, b)
        assert _same(f.streaming_topk(*a), o.streaming_topk(*a))
    bk = f.backup_tppr()
    assert f._sched is None and sp2.rolled_back and sp2.hits == 3
    _state_equal(f, o, probe)
    f.restore_tppr(bk)
    a = _args(st, neg, bs, 18)
    assert _same(f.streaming_topk(*a), o.streaming_topk(*a))
    f.sync_check()
    assert f.counters()["timeouts"] == 0


class _Model:
    """The part of the reference's TGN that eval_edge_prediction touches: compute_edge_probabilities queries the
    streaming finder with [src | dst | neg] exactly as model/tgn_model.py:79-83 + embedding_module.py:217-225 do."""

    def __init__(self, finder):
        self.embedding_module = types.SimpleNamespace(tppr_strategy="streaming", tppr_finder=finder)
        self.digests = []

    def eval(self):
        return self

    def compute_edge_probabilities(self, src, dst, neg, ts, eidx, n_neighbors, train=False):
        import torch
        nodes = np.concatenate([src, dst, neg]).astype(np.int32)
        buf, rows = self.embedding_module.tppr_finder.streaming_topk_raw(nodes, np.concatenate([ts, ts, ts]), eidx)
        self.digests.append(buf.cpu().numpy().tobytes())
        n = len(src)
        w = buf[rows * 20 * 12: rows * 20 * 16].view(torch.float32).view(rows, 20).sum(1)      # instance 0 weights
        return torch.sigmoid(w[:n] - w[2 * n:]).reshape(-1, 1), torch.sigmoid(w[2 * n:] - w[n:2 * n]).reshape(-1, 1)


def test_eval_edge_prediction_with_and_without_the_schedule():
    from zebra_b200 import metrics as M
    from zebra_b200.batches import RandEdgeSampler
    st = synth.make_stream("wikipedia", n_edges=5_000)
    N = st.shape.table_nodes
    lo, hi = 3_000, 4_777
    data = types.SimpleNamespace(sources=st.src[lo:hi].astype(np.int64), destinations=st.dst[lo:hi].astype(np.int64),
                                 timestamps=st.ts[lo:hi], edge_idxs=st.eidx[lo:hi], n_interactions=hi - lo)
    res = {}
    for on in (True, False):
        f = _finder(N)
        f.compute_val_tppr(st.src[:lo], st.dst[:lo], st.ts[:lo], st.eidx[:lo])
        m = _Model(f)
        M.PASS_SCHEDULE = on
        try:
            out = M.eval_edge_prediction(m, RandEdgeSampler(st.src, st.dst, seed=2), data, 20, 200)
        finally:
            M.PASS_SCHEDULE = True
        res[on] = (out, m.digests, M.LAST_PASS)
    assert res[True][0] == res[False][0]
    assert res[True][1] == res[False][1] and len(res[True][1]) == 9
    assert res[True][2] == {"batches": 9, "served": 9, "launches": 1, "rolled_back": False}
    assert res[False][2] is None
