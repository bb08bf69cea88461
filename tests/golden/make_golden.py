"""Generate golden fixtures by EXECUTING THE REFERENCE ITSELF (build container only).

Imports /root/reference/utils/util.py (numba jitclass `tppr_finder`, `NeighborFinder`) unmodified,
feeds it seeded synthetic streams and records inputs + outputs under tests/golden/*.npz.  The
reference has no tests or golden vectors of its own, so these recordings are what pins the
oracle (tests/test_oracle.py) and, through it, the CUDA path.

Run:  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py
The GPU box has no /root/reference; nothing at test time imports it — only these fixtures travel.
"""
from __future__ import annotations

import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

from zebra_b200 import synth  # noqa: E402


def _ref():
    sys.path.insert(0, "/root/reference")
    from utils import util as ref_util  # noqa
    return ref_util


def digest(arrs):
    h = hashlib.sha256()
    for a in arrs:
        h.update(np.ascontiguousarray(a).tobytes())
    return np.frombuffer(h.digest(), dtype=np.uint8).copy()


def dump_state(f, nodes):
    """Reference state for `nodes`: per instance (counts, eidx, node, ts, score, norm) padded to k."""
    k, n_t = f.k, f.n_tppr
    cnt = np.zeros((n_t, len(nodes)), np.int32)
    e = np.zeros((n_t, len(nodes), k), np.int64); nd = np.zeros_like(e)
    t = np.zeros((n_t, len(nodes), k), np.float64); s = np.zeros_like(t)
    nm = np.zeros((n_t, len(nodes)), np.float64)
    for j in range(n_t):
        for a, v in enumerate(nodes):
            d = f.PPR_list[j][int(v)]
            nm[j, a] = f.norm_list[j][int(v)]
            for b, (key, val) in enumerate(d.items()):
                e[j, a, b], nd[j, a, b], t[j, a, b], s[j, a, b] = key[0], key[1], key[2], val
            cnt[j, a] = len(d)
    return dict(st_cnt=cnt, st_eidx=e, st_node=nd, st_ts=t, st_score=s, st_norm=nm)


def run_stream_case(ref, name, src, dst, neg, ts, eidx, num_nodes, k, alphas, betas, bs, keep_batches=2,
                    no_fake=False):
    import numba as nb
    f = ref.tppr_finder(num_nodes, k, len(alphas), list(alphas), list(betas))
    E = len(src)
    digests, kept = [], {}
    nb_ = (E + bs - 1) // bs
    for b in range(nb_):
        lo, hi = b * bs, min(E, (b + 1) * bs)
        if no_fake:
            nodes = np.concatenate([src[lo:hi], dst[lo:hi]]).astype(np.int32)
            tt = np.concatenate([ts[lo:hi]] * 2)
            out = f.streaming_topk_no_fake(nodes, tt, eidx[lo:hi].astype(np.int64))
        else:
            nodes = np.concatenate([src[lo:hi], dst[lo:hi], neg[lo:hi]]).astype(np.int32)
            tt = np.concatenate([ts[lo:hi]] * 3)
            out = f.streaming_topk(nodes, tt, eidx[lo:hi].astype(np.int64))
        flat = [a for lst in out for a in lst]
        digests.append(digest(flat))
        if b >= nb_ - keep_batches:
            for q, nm in enumerate(("o_node", "o_eidx", "o_dt", "o_w")):
                kept[f"{nm}_{b}"] = np.stack(out[q])
    touched = np.unique(np.concatenate([src, dst]))
    rng = np.random.default_rng(3)
    sample = touched if len(touched) <= 400 else np.sort(rng.choice(touched, 400, replace=False))
    st = dump_state(f, sample)
    path = os.path.join(HERE, f"{name}.npz")
    np.savez_compressed(path, src=src, dst=dst, neg=neg if neg is not None else np.zeros(0, np.int32), ts=ts,
                        eidx=eidx, num_nodes=num_nodes, k=k, alphas=np.array(alphas), betas=np.array(betas),
                        bs=bs, no_fake=int(no_fake), digests=np.stack(digests), state_nodes=sample, **st, **kept)
    print(name, "edges", E, "->", os.path.getsize(path) // 1024, "KiB")
    return f


def adversarial(seed, n_nodes, E, p_loop=0.1, p_dup_ts=0.5, replay=False):
    rng = np.random.default_rng(seed)
    src = rng.integers(1, n_nodes + 1, E).astype(np.int32)
    dst = rng.integers(1, n_nodes + 1, E).astype(np.int32)
    loop = rng.random(E) < p_loop
    dst = np.where(loop, src, dst).astype(np.int32)
    rep = rng.random(E) < 0.3
    rep[0] = False
    for i in range(1, E):
        if rep[i]:
            src[i], dst[i] = (src[i - 1], dst[i - 1]) if rng.random() < 0.5 else (dst[i - 1], src[i - 1])
    gaps = np.where(rng.random(E) < p_dup_ts, 0.0, rng.random(E) * 3.7)
    ts = 1.2e6 + np.cumsum(gaps)
    eidx = np.arange(1, E + 1, dtype=np.int64)
    if replay:  # replayed edge ids with identical (src,dst,ts): the polluted-eval-table case (H9)
        h = E // 2
        src[h:], dst[h:], ts[h:], eidx[h:] = src[:E - h], dst[:E - h], ts[:E - h], eidx[:E - h]
    neg = rng.integers(1, n_nodes + 1, E).astype(np.int32)
    return src, dst, neg, ts, eidx


def main():
    ref = _ref()
    import numba as nb

    # 1. Wikipedia-shaped prefix, the reference's own CPU-runnable configuration (configs[0])
    st = synth.make_stream("wikipedia", n_edges=6000)
    neg = synth.make_negatives(st, 200)
    run_stream_case(ref, "stream_wikipedia6k", st.src, st.dst, neg, st.ts, st.eidx, st.shape.table_nodes, 20,
                    [0.1, 0.1], [0.5, 0.95], 200)
    # 2. unipartite with self-loops and replies
    st = synth.make_stream("superuser", n_edges=5000,
                           shape=synth.StreamShape("su_small", 3000, 0, 5000, 0, 30.0, 0.05, 0.3, 0.01))
    neg = synth.make_negatives(st, 200)
    run_stream_case(ref, "stream_unipartite5k", st.src, st.dst, neg, st.ts, st.eidx, 3001, 20,
                    [0.1, 0.1], [0.5, 0.95], 200)
    # 3. adversarial: tiny id space, self-loops, repeated pairs, duplicate fractional timestamps
    a = adversarial(5, 12, 1500)
    run_stream_case(ref, "stream_adv_k3", *a, 13, 3, [0.1, 0.0, 0.3], [0.5, 0.95, 0.7], 50, keep_batches=3)
    a = adversarial(6, 40, 2000)
    run_stream_case(ref, "stream_adv_k20", *a, 41, 20, [0.1, 0.1], [0.5, 0.95], 200)
    a = adversarial(7, 25, 1200, replay=True)
    run_stream_case(ref, "stream_adv_replay", *a, 26, 5, [0.1, 0.2], [0.5, 0.9], 100)
    # 4. exact ties: alpha=beta=0.5 makes many equal scores
    a = adversarial(8, 10, 800, p_loop=0.05)
    run_stream_case(ref, "stream_ties", *a, 11, 2, [0.5, 0.5], [0.5, 0.25], 40)
    a = adversarial(9, 30, 1500, p_loop=0.0)
    run_stream_case(ref, "stream_ties_k8", *a, 31, 8, [0.5], [0.5], 100)
    # 5. no-fake variant, k=1, k=31
    a = adversarial(10, 50, 1000)
    run_stream_case(ref, "stream_nofake", a[0], a[1], None, a[3], a[4], 51, 7, [0.1, 0.1], [0.5, 0.95], 100,
                    no_fake=True)
    a = adversarial(11, 30, 600)
    run_stream_case(ref, "stream_k1", *a, 31, 1, [0.1], [0.5], 100)
    a = adversarial(12, 200, 3000, p_loop=0.02)
    run_stream_case(ref, "stream_k31", *a, 201, 31, [0.15, 0.05], [0.6, 0.9], 200)

    # 6. protocol: fill (compute_val_tppr) / backup / restore / restore_val aliasing + single_streaming_topk
    st = synth.make_stream("wikipedia", n_edges=3000, seed=77)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes
    f = ref.tppr_finder(N, 10, 2, [0.1, 0.1], [0.5, 0.95])

    def batch(lo, hi):
        nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
        return nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi].astype(np.int64)

    outs = {}
    f.streaming_topk(*batch(0, 200))                                   # train
    f.reset_tppr()
    f.compute_val_tppr(st.src[:1000].astype(np.int64), st.dst[:1000].astype(np.int64), st.ts[:1000],
                       st.eidx[:1000].astype(np.int64))                # fill (epoch 1)
    bk = f.backup_tppr()
    outs["val1"] = np.stack([np.stack(x) for x in f.streaming_topk(*batch(1000, 1200))])
    bk2 = f.backup_tppr()
    f.restore_tppr(bk)
    outs["nn_val1"] = np.stack([np.stack(x) for x in f.streaming_topk(*batch(1000, 1200))])
    f.restore_tppr(bk2)
    f.reset_tppr()                                                     # epoch 2 train
    f.streaming_topk(*batch(0, 200))
    f.reset_tppr()
    f.restore_val_tppr()                                               # epoch 2 fill
    outs["val2"] = np.stack([np.stack(x) for x in f.streaming_topk(*batch(1000, 1200))])
    s0 = f.single_streaming_topk(*batch(1200, 1400), 1)
    outs["single"] = np.stack([np.asarray(x, dtype=np.float64) for x in s0])
    sample = np.unique(np.concatenate([st.src[:1400], st.dst[:1400]]))[:300]
    np.savez_compressed(os.path.join(HERE, "protocol_alias.npz"), src=st.src, dst=st.dst, neg=neg, ts=st.ts,
                        eidx=st.eidx, num_nodes=N, k=10, alphas=np.array([0.1, 0.1]), betas=np.array([0.5, 0.95]),
                        state_nodes=sample, **dump_state(f, sample), **outs)
    print("protocol_alias done")

    # 7. pruning strategy (configs[2], MOOC-shaped prefix)
    st = synth.make_stream("mooc", n_edges=8000)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes

    class D:  # the four attributes get_neighbor_finder reads (util.py:90-96)
        sources, destinations, edge_idxs, timestamps = st.src[:6000], st.dst[:6000], st.eidx[:6000], st.ts[:6000]
    nf = ref.get_neighbor_finder(D)
    cases = [(10, 2, 0.1, 0.5, 20), (10, 2, 0.1, 0.95, 20), (5, 3, 0.0, 0.7, 8), (20, 1, 0.2, 0.5, 20), (3, 2, 0.1, 0.5, 20)]
    lo, hi = 5400, 6000
    q_nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
    q_ts = np.concatenate([st.ts[lo:hi]] * 3)
    # also query beyond every edge time and at a node with no history
    q_nodes = np.concatenate([q_nodes, np.array([1, 2, N - 1, int(st.dst[0])], np.int32)])
    q_ts = np.concatenate([q_ts, np.array([1e9, 0.0, 1e9, st.ts[0]])])
    pr = {}
    for ci, (w, d, al, be, k) in enumerate(cases):
        o = [np.zeros((len(q_nodes), k), np.int32), np.zeros((len(q_nodes), k), np.int32),
             np.zeros((len(q_nodes), k), np.float32), np.zeros((len(q_nodes), k), np.float32)]
        nf.get_pruned_topk(q_nodes, q_ts, w, d, al, be, k, *o)
        for nm, a in zip(("node", "eidx", "dt", "w"), o):
            pr[f"c{ci}_{nm}"] = a
    np.savez_compressed(os.path.join(HERE, "pruning_mooc.npz"), src=st.src[:6000], dst=st.dst[:6000],
                        ts=st.ts[:6000], eidx=st.eidx[:6000], num_nodes=N, q_nodes=q_nodes, q_ts=q_ts,
                        cases=np.array(cases, dtype=np.float64), **pr)
    print("pruning_mooc done")

    # 8. numba argsort behaviour on tied inputs (pins oracle.numba_argsort)
    @nb.njit
    def asort(v):
        return np.argsort(v)
    rng = np.random.default_rng(21)
    vals, perms = [], []
    for n in (2, 5, 14, 15, 16, 21, 41, 41, 41, 63, 100):
        for levels in (2, 3, 7, 1000):
            v = rng.integers(0, levels, n).astype(np.float64) / 7.0
            p = asort(v)
            vv = np.full(100, np.nan); vv[:n] = v
            pp = np.full(100, -1, np.int64); pp[:n] = p
            vals.append(vv); perms.append(pp)
    np.savez_compressed(os.path.join(HERE, "numba_argsort.npz"), vals=np.stack(vals), perms=np.stack(perms))
    print("numba_argsort done")


if __name__ == "__main__":
    main()
