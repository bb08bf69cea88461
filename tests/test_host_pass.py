"""CPU tests of the host dependency pass behind zebra_tppr_plan_create (zebra_b200/csrc/tppr_api.cu:
fill_soa for one GPU, shard_work_list + fill_soa_sharded for node-sharded handles), through the host-only
test hook zebra_debug_host_pass.  The expectation is an independent pure-Python restatement of the rules
the span kernel relies on (tppr_device.cuh, "dependencies"):

  prev_x   span index of the previous edge that WROTE endpoint x (src/dst of an edge are written, a
           negative is only read), -1 if none;
  rthr_x   how many negative-sample reads of x's current version precede this write;
  signal   this edge's negative read is one of those counted reads (a later edge overwrites the node);
  dst1st   both endpoints are rewritten later and dst strictly sooner than src.
On a node-sharded handle only owned nodes (id & (world-1) == rank) are tracked and only the edges that
touch an owned node are staged, in span order."""
import ctypes as C

import numpy as np
import pytest

from zebra_b200 import _lib, synth


def _host_pass(N, world, rank, src, dst, neg, ts, eidx, bs):
    L = _lib.lib()
    n = len(src)
    fields = np.zeros((10, n), np.int32)
    ots = np.zeros(n, np.float64)
    lst = np.zeros(n, np.int32)
    ofs = np.zeros((n + bs - 1) // bs + 1, np.int64)
    cnt = C.c_int64(0)
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    rc = L.zebra_debug_host_pass(N, world, rank, p(src), p(dst), p(neg), p(ts), p(eidx), n, bs, p(fields), p(ots), p(lst), p(ofs),
                                 C.byref(cnt))
    assert rc == 0, L.zebra_last_error()
    c = cnt.value
    # the hook packs the ten fields back to back with the real count as row length
    return fields.reshape(-1)[:10 * c].reshape(10, c), ots[:c], lst[:c], ofs


def _expected(world, rank, src, dst, neg, bs):
    pm = world - 1
    own = lambda v: (int(v) & pm) == rank
    lastw, readers, recs, ofs = {}, {}, [], []
    n = len(src)
    for i in range(n):
        if i % bs == 0:
            ofs.append(len(recs))
        a, b = int(src[i]), int(dst[i])
        g = int(neg[i]) if neg is not None else a
        oa, ob = own(a), own(b)
        og = neg is not None and g != a and g != b and own(g)
        if not (oa or ob or og):
            continue
        q = len(recs)
        r = dict(i=i, src=a, dst=b, neg=g, prev_s=-1, prev_d=-1, prev_n=-1, rthr_s=0, rthr_d=0, signal=0, dst1st=0, oa=oa, ob=ob)
        if og:
            r["prev_n"] = lastw.get(g, -1)
            readers.setdefault(g, []).append(q)
        for node, o, pk, rk in ((a, oa, "prev_s", "rthr_s"), (b, ob and b != a, "prev_d", "rthr_d")):
            if o:
                r[pk] = lastw.get(node, -1)
                rd = readers.pop(node, [])
                r[rk] = len(rd)
                for x in rd:
                    recs[x]["signal"] = 1
                lastw[node] = i
        recs.append(r)
    ofs.append(len(recs))
    nextw = {}
    for r in reversed(recs):
        a, b = r["src"], r["dst"]
        if r["oa"] and r["ob"] and a != b and nextw.get(b, 1 << 62) < nextw.get(a, 1 << 62):
            r["dst1st"] = 1
        if r["oa"]:
            nextw[a] = r["i"]
        if r["ob"]:
            nextw[b] = r["i"]
    return recs, ofs


def _adversarial(n, N, seed, with_neg=True):
    rng = np.random.default_rng(seed)
    src = rng.integers(0, N, n).astype(np.int32)
    dst = rng.integers(0, N, n).astype(np.int32)
    neg = rng.integers(0, N, n).astype(np.int32)
    loops = rng.random(n) < 0.03
    dst[loops] = src[loops]
    neg[rng.random(n) < 0.05] = 0
    m = rng.random(n) < 0.05
    neg[m] = src[m]
    m = rng.random(n) < 0.05
    neg[m] = dst[m]
    for i in range(1, n):
        u = rng.random()
        if u < 0.08:                                 # the same pair again
            src[i], dst[i] = src[i - 1], dst[i - 1]
        elif u < 0.16:                               # the reversed pair: both endpoints rewritten by one edge
            src[i], dst[i] = dst[i - 1], src[i - 1]
    ts = np.cumsum(rng.integers(0, 3, n)).astype(np.float64)
    eidx = np.arange(1, n + 1, dtype=np.int64)
    return src, dst, (neg if with_neg else None), ts, eidx


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("N,n,bs,with_neg", [(5, 700, 7, True), (37, 1500, 200, True), (37, 1500, 200, False), (1000, 4001, 333, True)])
def test_host_pass_matches_restatement(world, N, n, bs, with_neg):
    src, dst, neg, ts, eidx = _adversarial(n, N, 100 * world + N + n, with_neg)
    for rank in range(world):
        f, ots, lst, ofs = _host_pass(N, world, rank, src, dst, neg, ts, eidx, bs)
        recs, eofs = _expected(world, rank, src, dst, neg, bs)
        assert len(lst) == len(recs) and np.array_equal(lst, [r["i"] for r in recs]), f"rank {rank}: work list"
        assert np.array_equal(ofs, eofs), f"rank {rank}: batch offsets"
        assert np.array_equal(f[0], src[lst]) and np.array_equal(f[1], dst[lst]) and np.array_equal(f[3], eidx[lst])
        assert np.array_equal(f[2], (neg if neg is not None else src)[lst])
        assert np.array_equal(ots, ts[lst])
        for name, row in (("prev_s", 4), ("prev_d", 5), ("prev_n", 6), ("rthr_s", 7), ("rthr_d", 8)):
            want = np.array([r[name] for r in recs], np.int32)
            bad = np.where(f[row] != want)[0]
            assert len(bad) == 0, f"rank {rank}/{world} {name}: record {bad[:5]} got {f[row][bad[:5]]} want {want[bad[:5]]}"
        assert np.array_equal(f[9] & 1, [r["signal"] for r in recs]), f"rank {rank}: signal flags"
        assert np.array_equal((f[9] >> 1) & 1, [r["dst1st"] for r in recs]), f"rank {rank}: dst-first flags"
        assert not (f[9] >> 2).any()


def test_host_pass_on_a_hub_stream_and_every_edge_is_staged_somewhere():
    st = synth.make_stream("wikipedia", n_edges=20000, seed=3)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes
    src, dst, neg = st.src.astype(np.int32), st.dst.astype(np.int32), neg.astype(np.int32)
    staged = np.zeros(st.n_edges, np.int32)
    for rank in range(4):
        f, _, lst, ofs = _host_pass(N, 4, rank, src, dst, neg, st.ts, st.eidx.astype(np.int64), 200)
        recs, eofs = _expected(4, rank, src, dst, neg, 200)
        assert np.array_equal(lst, [r["i"] for r in recs]) and np.array_equal(ofs, eofs)
        for name, row in (("prev_s", 4), ("prev_d", 5), ("prev_n", 6), ("rthr_s", 7), ("rthr_d", 8)):
            assert np.array_equal(f[row], [r[name] for r in recs]), name
        assert np.array_equal(f[9] & 3, [r["signal"] | (r["dst1st"] << 1) for r in recs])   # (bits 2.. mark chain endpoints)
        staged[lst] += 1
    assert staged.min() >= 1 and staged.max() <= 3          # every edge has an owner; at most src, dst and negative owners


def test_host_pass_rejects_bad_input():
    L = _lib.lib()
    src = np.array([1, 2, 9], np.int32)
    dst = np.array([2, 3, 1], np.int32)
    ts = np.zeros(3)
    eidx = np.arange(3, dtype=np.int64)
    out = np.zeros((10, 3), np.int32)
    o2, o3, o4 = np.zeros(3), np.zeros(3, np.int32), np.zeros(3, np.int64)
    cnt = C.c_int64(0)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for world in (1, 2):
        rc = L.zebra_debug_host_pass(5, world, 0, p(src), p(dst), None, p(ts), p(eidx), 3, 2, p(out), p(o2), p(o3), p(o4), C.byref(cnt))
        assert rc != 0 and b"out of range" in L.zebra_last_error()
    assert L.zebra_debug_host_pass(5, 3, 0, p(src), p(dst), None, p(ts), p(eidx), 3, 2, p(out), p(o2), p(o3), p(o4), C.byref(cnt)) != 0
