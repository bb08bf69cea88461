"""GPU parity of the pruning-strategy query (prune_topk_kernel through the C ABI and the NeighborFinder
drop-in) against recordings of the reference (tests/golden/pruning_mooc.npz) and against the C oracle."""
import numpy as np
import pytest

import oracle
from tests import helpers as H
from zebra_b200 import synth

pytestmark = pytest.mark.gpu


def _nf(src, dst, eidx, ts, num_nodes):
    from zebra_b200.neighbor_finder import NeighborFinder
    return NeighborFinder.from_edges(src, dst, eidx, ts, num_nodes)


def _zeros(n, k):
    return [np.zeros((n, k), np.int32), np.zeros((n, k), np.int32), np.zeros((n, k), np.float32), np.zeros((n, k), np.float32)]


def test_golden_pruning_bit_exact():
    g = H.load("pruning_mooc")
    nf = _nf(g["src"], g["dst"], g["eidx"], g["ts"], int(g["num_nodes"]))
    q_nodes, q_ts = g["q_nodes"], g["q_ts"]
    for ci, (w, d, al, be, k) in enumerate(g["cases"]):
        o = _zeros(len(q_nodes), int(k))
        nf.get_pruned_topk(q_nodes, q_ts, int(w), int(d), float(al), float(be), int(k), *o)
        for nm, a in zip(("node", "eidx", "dt", "w"), o):
            assert np.array_equal(a, g[f"c{ci}_{nm}"]), (ci, nm)


@pytest.mark.parametrize("name,n_edges,width,depth,alpha,beta,k", [
    ("mooc", 60000, 10, 2, 0.1, 0.5, 20), ("mooc", 60000, 10, 2, 0.1, 0.95, 20),
    ("reddit", 40000, 10, 2, 0.1, 0.5, 20), ("wikipedia", 30000, 20, 2, 0.1, 0.5, 20),
    ("superuser", 50000, 5, 3, 0.1, 0.5, 10), ("mooc", 20000, 10, 2, 0.0, 0.5, 20),
    ("mooc", 20000, 3, 4, 0.2, 0.7, 32), ("mooc", 20000, 10, 1, 0.1, 0.5, 1), ("mooc", 20000, 40, 2, 0.1, 0.5, 5),
])
def test_seeded_pruning_vs_oracle(name, n_edges, width, depth, alpha, beta, k):
    st = synth.make_stream(name, n_edges=n_edges)
    N = st.shape.table_nodes
    split = int(0.7 * n_edges)
    csr = oracle.build_csr(st.src[:split], st.dst[:split], st.eidx[:split], st.ts[:split], N)
    ref = oracle.OracleNeighborFinder(*csr)
    nf = _nf(st.src[:split], st.dst[:split], st.eidx[:split], st.ts[:split], N)
    neg = synth.make_negatives(st, 200)
    for lo in (0, split // 2, split - 200, n_edges - 400):
        hi = lo + 200
        q = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        t = np.concatenate([st.ts[lo:hi]] * 3)
        a, b = _zeros(len(q), k), _zeros(len(q), k)
        nf.get_pruned_topk(q, t, width, depth, alpha, beta, k, *a)
        ref.get_pruned_topk(q, t, width, depth, alpha, beta, k, *b)
        for nm, x, y in zip(("node", "eidx", "dt", "w"), a, b):
            assert np.array_equal(x, y), (name, lo, nm)
        dev = nf.get_pruned_topk_device(q, t, width, depth, alpha, beta, k)
        for x, y in zip(dev, b):
            assert np.array_equal(x.cpu().numpy(), y)


def test_pruning_leaves_untouched_rows_and_rejects_bad_input():
    st = synth.make_stream("mooc", n_edges=5000)
    N = st.shape.table_nodes
    nf = _nf(st.src, st.dst, st.eidx, st.ts, N)
    ref = oracle.OracleNeighborFinder(*oracle.build_csr(st.src, st.dst, st.eidx, st.ts, N))
    # queries at time 0 have no earlier neighbour: the caller's rows must survive, as in the reference
    q = np.array([st.src[0], st.dst[0], 0, st.src[100]], np.int32)
    t = np.array([0.0, 0.0, 1e9, st.ts[-1] + 1])
    a = [np.full((4, 20), 7, dt) for dt in (np.int32, np.int32, np.float32, np.float32)]
    b = [x.copy() for x in a]
    nf.get_pruned_topk(q, t, 10, 2, 0.1, 0.5, 20, *a)
    ref.get_pruned_topk(q, t, 10, 2, 0.1, 0.5, 20, *b)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    assert (a[0][0] == 7).all() and (a[0][2] == 7).all() and not (a[0][3] == 7).all()
    with pytest.raises(RuntimeError):
        nf.get_pruned_topk(np.array([N + 5], np.int32), np.array([1.0]), 10, 2, 0.1, 0.5, 20, *_zeros(1, 20))
    with pytest.raises(RuntimeError):
        nf.get_pruned_topk(q, t, 10, 2, 0.1, 0.5, 40, *_zeros(4, 40))
    assert nf.launches() >= 1
    e, i, ts = nf.find_before(int(st.src[100]), st.ts[-1] + 1)
    assert len(e) == (np.sum(st.src == st.src[100]) + np.sum(st.dst == st.src[100]))
