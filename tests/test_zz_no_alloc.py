"""zebra_prune_topk_static — the allocation-free variant of the pruning query (static device work area) —
through the C ABI against zebra_prune_topk on the same inputs: bit-exact, the tolerance of every prune test here.
Collected last on purpose.  No counterpart in the reference."""
import ctypes as C

import numpy as np
import pytest

from zebra_b200 import synth

pytestmark = pytest.mark.gpu


def _setup(n_edges=30000, name="mooc"):
    import torch
    from zebra_b200._lib import lib
    from zebra_b200.neighbor_finder import build_csr
    st = synth.make_stream(name, n_edges=n_edges)
    N = st.shape.table_nodes
    indptr, adj_node, adj_eidx, adj_ts = build_csr(st.src, st.dst, st.eidx, st.ts, N)
    L = lib()
    h = C.c_void_p()
    rc = L.zebra_csr_create(C.byref(h), indptr.ctypes.data_as(C.c_void_p), adj_node.ctypes.data_as(C.c_void_p),
                            adj_eidx.ctypes.data_as(C.c_void_p), adj_ts.ctypes.data_as(C.c_void_p), N, len(adj_node), 0)
    assert rc == 0, L.zebra_last_error().decode()
    dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (indptr, adj_node, adj_eidx, adj_ts)]
    return st, N, L, h, dev


def _queries(st, lo, n):
    q = np.concatenate([st.src[lo:lo + n], st.dst[lo:lo + n]]).astype(np.int32)
    t = np.concatenate([st.ts[lo:lo + n]] * 2).astype(np.float64)
    return np.ascontiguousarray(q), np.ascontiguousarray(t)


def _existing(L, h, q, t, width, depth, alpha, beta, k):
    nb = L.zebra_prune_out_bytes(len(q), k)
    out = np.zeros(nb, np.uint8)
    rc = L.zebra_prune_topk(h, q.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p), len(q), width, depth,
                            alpha, beta, k, None, out.ctypes.data_as(C.c_void_p), None)
    assert rc == 0, L.zebra_last_error().decode()
    return out


def _static(L, dev, N, q, t, width, depth, alpha, beta, k, poison=0):
    import torch
    nb = L.zebra_prune_out_bytes(len(q), k)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    out = torch.full((nb,), poison, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free0, _ = torch.cuda.mem_get_info()
    rc = L.zebra_prune_topk_static(*(C.c_void_p(a.data_ptr()) for a in dev), N, C.c_void_p(dq.data_ptr()),
                                   C.c_void_p(dt.data_ptr()), len(q), width, depth, alpha, beta, k,
                                   C.c_void_p(out.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, L.zebra_last_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    free1, _ = torch.cuda.mem_get_info()
    return out.cpu().numpy(), (free0, free1)


def test_static_entry_point_matches_existing_on_two_calls_in_a_row():
    st, N, L, h, dev = _setup()
    try:
        # first call: many queries, wide; second call right after it: fewer, other parameters (finds the first
        # call's states in the shared work area); third: the first again
        cases = [(_queries(st, 20000, 1500), (20, 2, 0.1, 0.5, 20)),
                 (_queries(st, 9000, 300), (3, 4, 0.2, 0.7, 32)),
                 (_queries(st, 20000, 1500), (20, 2, 0.1, 0.5, 20))]
        for i, ((q, t), prm) in enumerate(cases):
            want = _existing(L, h, q, t, *prm)
            got, (free0, free1) = _static(L, dev, N, q, t, *prm, poison=0xAB if i else 0)
            assert got.shape == want.shape
            assert np.array_equal(got, want), (i, int((got != want).sum()))
            if i:       # the device's free memory did not shrink across the call (the first call may load the module)
                assert free1 >= free0, (i, free0, free1)
    finally:
        L.zebra_csr_destroy(h)


def test_static_entry_point_limits_and_bad_ids():
    import torch
    st, N, L, h, dev = _setup(n_edges=5000)
    try:
        q, t = _queries(st, 3000, 50)
        q = q.copy()
        q[3], q[7] = N + 11, -2                       # ids the host cannot check: neutralised, flagged -1
        got, _ = _static(L, dev, N, q, t, 10, 2, 0.1, 0.5, 20)
        k, nq = 20, len(q)
        filled = got[16 * nq * k:].view(np.int32)
        assert filled[3] == -1 and filled[7] == -1 and (np.delete(filled, [3, 7]) >= 0).all()
        good = np.ones(nq, bool)
        good[[3, 7]] = False
        want = _existing(L, h, np.ascontiguousarray(q[good]), np.ascontiguousarray(t[good]), 10, 2, 0.1, 0.5, 20)
        ng = int(good.sum())
        for pl in range(4):
            a = got[pl * 4 * nq * k:(pl + 1) * 4 * nq * k].view(np.int32).reshape(nq, k)
            b = want[pl * 4 * ng * k:(pl + 1) * 4 * ng * k].view(np.int32).reshape(ng, k)
            assert np.array_equal(a[good], b)
            assert (a[~good] == 0).all()
        # beyond the limits: a clear error, nothing launched
        dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
        out = torch.zeros(L.zebra_prune_out_bytes(70000, 20), dtype=torch.uint8, device="cuda")
        args = [C.c_void_p(a.data_ptr()) for a in dev] + [N, C.c_void_p(dq.data_ptr()), C.c_void_p(dt.data_ptr())]
        assert L.zebra_prune_topk_static(*args, 70000, 10, 2, 0.1, 0.5, 20, C.c_void_p(out.data_ptr()), None) != 0
        assert b"65536" in L.zebra_last_error()
        assert L.zebra_prune_topk_static(*args, nq, 1000, 2, 0.1, 0.5, 20, C.c_void_p(out.data_ptr()), None) != 0
        assert b"static work area" in L.zebra_last_error()
        assert L.zebra_prune_topk_static(*args, nq, 10, 2, 0.1, 0.5, 33, C.c_void_p(out.data_ptr()), None) != 0
        assert L.zebra_prune_topk_static(*args, nq, 10, 2, 0.1, 0.5, 20, None, None) != 0
        torch.cuda.synchronize()
    finally:
        L.zebra_csr_destroy(h)
