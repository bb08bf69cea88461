"""Shared helpers for replaying golden fixtures through a finder-like object."""
import hashlib
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
STREAM_CASES = ["stream_wikipedia6k", "stream_unipartite5k", "stream_adv_k3", "stream_adv_k20", "stream_adv_replay",
                "stream_ties", "stream_ties_k8", "stream_nofake", "stream_k1", "stream_k31",
                # tests/golden/make_golden_late.py: timestamps of ~2e8 s, where float32(ts) is on a 16 s grid and the
                # reference's time differences are inexact and sometimes negative; a million-node id space
                "stream_su_late", "stream_wt_late"]
STREAM_CASES_LATE = ["stream_su_late", "stream_wt_late"]


def load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def digest(arrs):
    h = hashlib.sha256()
    for a in arrs:
        h.update(np.ascontiguousarray(a).tobytes())
    return np.frombuffer(h.digest(), dtype=np.uint8).copy()


def batches(g):
    """Yield (b, nodes int32, ts f64, eidx int64) exactly as make_golden.py fed the reference."""
    src, dst, neg, ts, eidx = g["src"], g["dst"], g["neg"], g["ts"], g["eidx"]
    bs, E, no_fake = int(g["bs"]), len(g["src"]), bool(g["no_fake"])
    for b in range((E + bs - 1) // bs):
        lo, hi = b * bs, min(E, (b + 1) * bs)
        parts = [src[lo:hi], dst[lo:hi]] + ([] if no_fake else [neg[lo:hi]])
        yield b, np.concatenate(parts).astype(np.int32), np.concatenate([ts[lo:hi]] * len(parts)), eidx[lo:hi].astype(np.int64)


def replay(finder, g):
    """Run all batches; return per-batch digests and the kept full outputs of the last batches."""
    no_fake = bool(g["no_fake"])
    digs, outs = [], {}
    for b, nodes, tt, ee in batches(g):
        out = finder.streaming_topk_no_fake(nodes, tt, ee) if no_fake else finder.streaming_topk(nodes, tt, ee)
        digs.append(digest([a for lst in out for a in lst]))
        outs[b] = [np.stack(x) for x in out]
    return np.stack(digs), outs


EMBED_CASES = [("streaming", "wikipedia", 2400, 172), ("pruning", "mooc", 2400, 4), ("streaming", "su_small", 1200, 1)]


def embedding_inputs(name, n_edges, d_e):
    """Seeded inputs shared by tests/golden/make_golden_embedding.py (which feeds them to the reference
    module) and the GPU parity test: stream, negatives, edge features [E+1,d_e], memory [N,100]."""
    from zebra_b200 import synth
    if name == "su_small":
        st = synth.make_stream("superuser", n_edges=n_edges,
                               shape=synth.StreamShape("su_small", 3000, 0, n_edges, 0, 30.0, 0.05, 0.3, 0.01))
    else:
        st = synth.make_stream(name, n_edges=n_edges)
    neg = synth.make_negatives(st, 200)
    N = st.shape.table_nodes
    rng = np.random.default_rng(17)
    ef = rng.standard_normal((n_edges + 1, d_e)).astype(np.float32)
    ef[0] = 0
    if d_e == 1:
        ef[:] = 0                                   # the reference's zeros((E+1,1)) substitute (train.py:133-136)
    memory = (rng.standard_normal((N, 100)) * 0.5).astype(np.float32)
    return st, neg, N, ef, memory
