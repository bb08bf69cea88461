"""The prune's exact-tie ordering in isolation (tppr_device.cuh tie_order: numba's quicksort partition
phase replayed in registers, two ranges per round) against the oracle's restatement of numba's argsort
(oracle/tppr_oracle.c zo_numba_argsort, itself pinned on recordings of numba 0.65: tests/golden/
numba_argsort.npz).  Reference call site: utils/util.py:555."""
import ctypes as C

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


def _ranks(values, ns, stride=64):
    from zebra_b200._lib import check, lib
    v = np.ascontiguousarray(values, np.int32)
    n = np.ascontiguousarray(ns, np.int32)
    out = np.full_like(v, -1)
    check(lib().zebra_debug_tie_order(v.ctypes.data_as(C.c_void_p), n.ctypes.data_as(C.c_void_p), len(n), stride,
                                      out.ctypes.data_as(C.c_void_p)))
    return out


def _cases(rng, count):
    vals, ns = np.zeros((count, 64), np.int32), np.zeros(count, np.int32)
    for c in range(count):
        n = int(rng.integers(2, 64))
        kind = c % 6
        if kind == 0:
            v = rng.integers(0, 3, n)                       # very few distinct values
        elif kind == 1:
            v = rng.integers(0, max(2, n // 3), n)
        elif kind == 2:                                       # two ascending runs + one element, as the prune feeds it
            a = int(rng.integers(1, n))
            v = np.concatenate([np.sort(rng.integers(0, 12, a)), np.sort(rng.integers(0, 12, n - a))])
            v[-1] = rng.integers(0, 12)
        elif kind == 3:
            v = np.sort(rng.integers(0, 8, n))[::-1]        # descending with ties
        elif kind == 4:
            v = np.zeros(n, np.int64)                         # all equal
            v[rng.integers(0, n, 3)] = rng.integers(0, 3, 3)
        else:
            v = rng.permutation(n) // 2                      # pairs
        vals[c, :n], ns[c] = v, n
    return vals, ns


def test_tie_order_matches_numba_argsort():
    rng = np.random.default_rng(5)
    vals, ns = _cases(rng, 6000)
    got = _ranks(vals, ns)
    bad = []
    for c in range(len(ns)):
        n = int(ns[c])
        order = oracle.numba_argsort(vals[c, :n].astype(np.float64))       # candidate at each sorted position
        want = np.empty(n, np.int32)
        want[order] = np.arange(n, dtype=np.int32)
        if not np.array_equal(got[c, :n], want):
            bad.append((c, n, vals[c, :n].tolist(), got[c, :n].tolist(), want.tolist()))
    assert not bad, f"{len(bad)} of {len(ns)} cases differ; first: {bad[0]}"


@pytest.mark.parametrize("name,n_edges", [("wikitalk", 60_000), ("reddit", 30_000)])
def test_stable_tie_mode_matches_oracle_stable_mode(name, n_edges):
    """tie_mode='stable' (zebra_tppr_set_tie_mode 1): equal scores keep insertion order — the oracle's
    tie_mode 1, bit for bit; and it differs from the exact mode only through exactly tied scores."""
    from zebra_b200 import synth
    from zebra_b200.finder import tppr_finder
    st = synth.make_stream(name, n_edges=n_edges)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes
    al, be = [0.1, 0.1], [0.5, 0.95]
    f = tppr_finder(N, 20, 2, al, be, tie_mode="stable")
    o = oracle.OracleFinder(N, 20, 2, al, be, tie_mode=1)
    plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, 200)
    out = plan.run(0, plan.n_batches, 32, out_slots=plan.n_batches)
    f.sync_check()
    for b in range(plan.n_batches):
        lo, hi = b * 200, min(n_edges, (b + 1) * 200)
        nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        ref = o.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
        got = plan.views(out, b, 0, plan.n_batches)
        for q in range(4):
            for j in range(2):
                assert np.array_equal(got[q][j].cpu().numpy(), ref[q][j]), (b, q, j)
    assert f.counters()["tie_prunes"] > 0
