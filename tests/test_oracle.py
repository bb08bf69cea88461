"""The oracle (oracle/tppr_oracle.c) against recordings of the reference itself (tests/golden)."""
import numpy as np
import pytest

import oracle
from tests import helpers as H


@pytest.mark.parametrize("name", H.STREAM_CASES)
def test_stream_bit_exact_vs_reference(name):
    g = H.load(name)
    f = oracle.OracleFinder(int(g["num_nodes"]), int(g["k"]), len(g["alphas"]), list(g["alphas"]), list(g["betas"]))
    digs, outs = H.replay(f, g)
    assert np.array_equal(digs, g["digests"]), f"first differing batch {np.argmax((digs != g['digests']).any(1))}"
    nb = len(digs)
    for key in g.files:
        if key.startswith("o_node_"):
            b = int(key.split("_")[-1])
            for q, nm in enumerate(("o_node", "o_eidx", "o_dt", "o_w")):
                assert np.array_equal(outs[b][q], g[f"{nm}_{b}"])
    # final state of sampled nodes: order, keys and fp64 scores bit-exact
    for j in range(f.n_tppr):
        for a, v in enumerate(g["state_nodes"]):
            e, n, t, s, nm = f.state(j, v)
            c = int(g["st_cnt"][j, a])
            assert len(e) == c
            assert np.array_equal(e, g["st_eidx"][j, a, :c]) and np.array_equal(n, g["st_node"][j, a, :c])
            assert np.array_equal(t, g["st_ts"][j, a, :c]) and np.array_equal(s, g["st_score"][j, a, :c])
            assert nm == g["st_norm"][j, a]


def test_numba_argsort_restatement():
    g = H.load("numba_argsort")
    for v, p in zip(g["vals"], g["perms"]):
        n = int(np.sum(~np.isnan(v)))
        assert np.array_equal(oracle.numba_argsort(v[:n]), p[:n])


def test_protocol_aliasing_matches_reference():
    g = H.load("protocol_alias")
    src, dst, neg, ts, eidx = g["src"], g["dst"], g["neg"], g["ts"], g["eidx"]
    f = oracle.OracleFinder(int(g["num_nodes"]), int(g["k"]), 2, list(g["alphas"]), list(g["betas"]))

    def batch(lo, hi):
        return (np.concatenate([src[lo:hi], dst[lo:hi], neg[lo:hi]]).astype(np.int32),
                np.concatenate([ts[lo:hi]] * 3), eidx[lo:hi].astype(np.int64))

    S = lambda o: np.stack([np.stack(x) for x in o])  # noqa: E731
    f.streaming_topk(*batch(0, 200))
    f.reset_tppr()
    f.compute_val_tppr(src[:1000].astype(np.int64), dst[:1000].astype(np.int64), ts[:1000], eidx[:1000].astype(np.int64))
    bk = f.backup_tppr()
    assert np.array_equal(S(f.streaming_topk(*batch(1000, 1200))).astype(np.float64), g["val1"])
    bk2 = f.backup_tppr()
    f.restore_tppr(bk)
    assert np.array_equal(S(f.streaming_topk(*batch(1000, 1200))).astype(np.float64), g["nn_val1"])
    f.restore_tppr(bk2)
    f.reset_tppr()
    f.streaming_topk(*batch(0, 200))
    f.reset_tppr()
    f.restore_val_tppr()
    assert np.array_equal(S(f.streaming_topk(*batch(1000, 1200))).astype(np.float64), g["val2"])
    s0 = f.single_streaming_topk(*batch(1200, 1400), 1)
    assert np.array_equal(np.stack([np.asarray(x, np.float64) for x in s0]), g["single"])
    for j in range(2):
        for a, v in enumerate(g["state_nodes"]):
            e, n, t, s, nm = f.state(j, v)
            c = int(g["st_cnt"][j, a])
            assert len(e) == c and np.array_equal(s, g["st_score"][j, a, :c]) and nm == g["st_norm"][j, a]
            assert np.array_equal(e, g["st_eidx"][j, a, :c])


def test_pruning_matches_reference():
    g = H.load("pruning_mooc")
    csr = oracle.build_csr(g["src"], g["dst"], g["eidx"], g["ts"], int(g["num_nodes"]))
    nf = oracle.OracleNeighborFinder(*csr)
    q_nodes, q_ts = g["q_nodes"], g["q_ts"]
    for ci, (w, d, al, be, k) in enumerate(g["cases"]):
        k = int(k)
        o = [np.zeros((len(q_nodes), k), np.int32), np.zeros((len(q_nodes), k), np.int32),
             np.zeros((len(q_nodes), k), np.float32), np.zeros((len(q_nodes), k), np.float32)]
        nf.get_pruned_topk(q_nodes, q_ts, int(w), int(d), float(al), float(be), k, *o)
        for nm, a in zip(("node", "eidx", "dt", "w"), o):
            assert np.array_equal(a, g[f"c{ci}_{nm}"]), (ci, nm)
