"""BASELINE.json configs at FULL size on the paths bench.py times — `plan.run` in windows of 64 batches
(the `value` leg) and `tppr_finder.stream_epoch` in windows of 128 batches (the `e2e` leg) — compared
row for row with the C oracle (reference: utils/util.py:447-469, 473-576), plus the final fp64 state
of >= 5,000 sampled nodes.  Bit-exact (np.array_equal) everywhere; no tolerance is used."""
import numpy as np
import pytest
import torch

import oracle
from zebra_b200 import synth

pytestmark = pytest.mark.gpu

K, BS = 20, 200
AL, BE = [0.1, 0.1], [0.5, 0.95]


def _finder(*a, **kw):
    from zebra_b200.finder import tppr_finder
    return tppr_finder(*a, **kw)


@pytest.mark.parametrize("name", ["wikipedia", "reddit", "mooc", "superuser", "wikitalk"])
def test_full_size_config_vs_oracle(name):
    st = synth.make_stream(name)
    E, N = st.n_edges, st.shape.table_nodes
    neg = synth.make_negatives(st, BS)
    f_plan, f_epoch = _finder(N, K, 2, AL, BE), _finder(N, K, 2, AL, BE)
    o = oracle.OracleFinder(N, K, 2, AL, BE)
    plan = f_plan.make_plan(st.src, st.dst, neg, st.ts, st.eidx, BS)
    nb, W = plan.n_batches, 64
    ring = plan.alloc_out(W)
    host = torch.empty(ring.numel(), dtype=torch.uint8).pin_memory()
    epoch = f_epoch.stream_epoch(st.src, st.dst, neg, st.ts, st.eidx, BS, window=128)
    for b0 in range(0, nb, W):
        n = min(W, nb - b0)
        plan.run(b0, n, W, out=ring, out_slots=W)
        host.copy_(ring)
        f_plan.sync_check()
        for b in range(b0, b0 + n):
            lo, hi = b * BS, min(E, (b + 1) * BS)
            nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
            ref = o.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
            got_p = plan.views(host, b, b0, W)
            got_e = next(epoch)
            for q in range(4):
                for j in range(2):
                    assert np.array_equal(got_p[q][j].numpy(), ref[q][j]), (name, "plan", b, q, j)
                    assert np.array_equal(got_e[q][j], ref[q][j]), (name, "epoch", b, q, j)
    assert next(epoch, None) is None
    rng = np.random.default_rng(3)
    deg = np.bincount(np.concatenate([st.src, st.dst]), minlength=N)
    sample = np.unique(np.concatenate([np.argsort(deg)[-500:], rng.choice(np.nonzero(deg)[0], min(6000, int((deg > 0).sum())), replace=False)]))
    assert len(sample) >= min(5000, int((deg > 0).sum()) // 2)
    for f in (f_plan, f_epoch):
        for j in range(2):
            for v, (e, n, t, s, nm) in zip(sample, f.state(j, sample)):
                oe, on, ot, os_, onm = o.state(j, v)
                assert np.array_equal(e, oe) and np.array_equal(n, on) and np.array_equal(t, ot)
                assert np.array_equal(s, os_), "fp64 scores must be bit-identical"
                assert nm == onm
        assert f.counters()["timeouts"] == 0


def test_abandoned_plan_leaves_no_stale_read_counts():
    """A plan abandoned part-way (its negative readers have signalled, the writers that consume the
    signals never ran) must not let the next run overwrite versions its own readers still need."""
    st = synth.make_stream("reddit", n_edges=20_000)
    N = st.shape.table_nodes
    neg = synth.make_negatives(st, BS)
    f = _finder(N, K, 2, AL, BE)
    plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, BS)
    out = plan.alloc_out(plan.n_batches)
    for cut in (1, 7, 33):
        plan.run(0, cut, 4, out=out, out_slots=plan.n_batches)      # abandoned after `cut` batches
        f.sync_check()
        plan2 = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, BS)
        f.reset_tppr()
        plan2.run(0, plan2.n_batches, 16, out=out, out_slots=plan2.n_batches)
        f.sync_check()
        o = oracle.OracleFinder(N, K, 2, AL, BE)
        for b in range(plan2.n_batches):
            lo, hi = b * BS, min(st.n_edges, (b + 1) * BS)
            nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
            ref = o.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
            got = plan2.views(out, b, 0, plan2.n_batches)
            for q in range(4):
                for j in range(2):
                    assert np.array_equal(got[q][j].cpu().numpy(), ref[q][j]), (cut, b, q, j)
        plan2.close()
        f.reset_tppr()
