"""Synthetic chronological edge streams shaped like the reference's datasets.

There is no network and no dataset on the box, so every measurement and parity
test runs on seeded synthetic streams whose node/edge counts, bipartite vs
unipartite structure, edge-feature width and time span follow the paper's
Table 3 (the shapes `BASELINE.json.configs` names).  The on-disk layout written
by `write_ml_format` is the one the reference loader reads
(/root/reference/utils/data_processing.py:68-81: `ml_{name}.csv` with columns
u,i,ts,label,idx and `ml_{name}.npy` with a zero row 0; ids are 1-based as in
/root/reference/utils/preprocess_data.py:46-55).

Nothing here touches the GPU; it is host-side input generation only.
"""
from __future__ import annotations

import dataclasses
import os

import numpy as np

DAY = 86400.0


@dataclasses.dataclass(frozen=True)
class StreamShape:
    name: str
    n_users: int          # bipartite: users; unipartite: all nodes
    n_items: int          # bipartite: items (ids offset by n_users); 0 = unipartite
    n_edges: int
    edge_dim: int         # 0 = no edge features (reference substitutes zeros((E+1,1)))
    span_days: float
    hub_share: float      # share of edges incident to the most popular item / node
    p_repeat: float       # bipartite: user repeats previous item; unipartite: reply to a recent edge
    self_loop: float = 0.0

    @property
    def n_nodes(self) -> int:
        return self.n_users + self.n_items

    @property
    def table_nodes(self) -> int:
        """Rows a T-PPR table needs: ids are 1-based, row 0 is the pad node."""
        return self.n_nodes + 1


# BASELINE.json configs[0..4]
SHAPES = {
    "wikipedia": StreamShape("wikipedia", 8227, 1000, 157_474, 172, 30.0, 0.01, 0.4),
    "reddit": StreamShape("reddit", 10_000, 984, 672_447, 172, 30.0, 0.08, 0.4),
    "mooc": StreamShape("mooc", 7047, 97, 411_749, 4, 30.0, 0.05, 0.3),
    "superuser": StreamShape("superuser", 194_085, 0, 1_443_339, 0, 2773.0, 0.015, 0.2, 0.001),
    "wikitalk": StreamShape("wikitalk", 1_140_149, 0, 7_833_140, 0, 2320.0, 0.015, 0.2, 0.001),
}
SEEDS = {"wikipedia": 1001, "reddit": 1002, "mooc": 1003, "superuser": 1004, "wikitalk": 1005}


@dataclasses.dataclass
class Stream:
    shape: StreamShape
    src: np.ndarray        # int32 [E], 1-based
    dst: np.ndarray        # int32 [E], 1-based
    ts: np.ndarray         # float64 [E], non-decreasing integer seconds
    eidx: np.ndarray       # int64 [E], 1..E
    dst_pool: np.ndarray   # unique destinations (negative sampling pool)

    @property
    def n_edges(self) -> int:
        return int(self.src.shape[0])


def _zipf_cdf(n: int, top_share: float) -> np.ndarray:
    """CDF of p(r) ∝ r^-a over ranks 1..n with a chosen so that p(1) = top_share."""
    if n <= 1:
        return np.ones(max(n, 1))
    ranks = np.arange(1, n + 1, dtype=np.float64)
    top_share = max(top_share, 1.0 / n)
    lo, hi = 0.0, 4.0
    for _ in range(60):
        a = 0.5 * (lo + hi)
        if 1.0 / np.sum(ranks ** -a) < top_share:
            lo = a
        else:
            hi = a
    w = ranks ** -(0.5 * (lo + hi))
    c = np.cumsum(w)
    return c / c[-1]


def _sample(cdf: np.ndarray, perm: np.ndarray, u: np.ndarray) -> np.ndarray:
    return perm[np.minimum(np.searchsorted(cdf, u, side="left"), len(cdf) - 1)]


def make_stream(name: str, seed: int | None = None, n_edges: int | None = None,
                shape: StreamShape | None = None) -> Stream:
    """Deterministic synthetic stream for one of SHAPES (optionally truncated to n_edges)."""
    shape = shape or SHAPES[name]
    seed = SEEDS.get(name, 1000) if seed is None else seed
    E = shape.n_edges if n_edges is None else int(n_edges)
    rng = np.random.default_rng(seed)

    if shape.n_items > 0:
        # bipartite: users mildly skewed, items with an explicit hub share
        ucdf = _zipf_cdf(shape.n_users, 4.0 / shape.n_users)
        icdf = _zipf_cdf(shape.n_items, shape.hub_share)
        uperm = rng.permutation(shape.n_users)
        iperm = rng.permutation(shape.n_items)
        user = _sample(ucdf, uperm, rng.random(E))
        item = _sample(icdf, iperm, rng.random(E))
        repeat = rng.random(E) < shape.p_repeat
        # a repeating edge takes the item of the user's previous non-repeating edge
        order = np.argsort(user, kind="stable")
        u_sorted = user[order]
        first = np.ones(E, dtype=bool)
        first[1:] = u_sorted[1:] != u_sorted[:-1]
        fresh = first | ~repeat[order]
        anchor = np.maximum.accumulate(np.where(fresh, np.arange(E), 0))
        item_sorted = item[order][anchor]
        item = np.empty_like(item)
        item[order] = item_sorted
        src = user + 1
        dst = item + 1 + shape.n_users
    else:
        N = shape.n_users
        scdf = _zipf_cdf(N, shape.hub_share)
        dcdf = _zipf_cdf(N, shape.hub_share * 0.5)
        sperm = rng.permutation(N)
        dperm = rng.permutation(N)
        s = _sample(scdf, sperm, rng.random(E))
        d = _sample(dcdf, dperm, rng.random(E))
        loop = rng.random(E) < shape.self_loop
        d = np.where(loop, s, d)
        # replies reverse the most recent base edge at lag 1..8
        reply = rng.random(E) < shape.p_repeat
        reply[:8] = False
        lag = rng.integers(1, 9, size=E)
        base_idx = np.maximum.accumulate(np.where(~reply, np.arange(E), 0))
        ref = base_idx[np.maximum(np.arange(E) - lag, 0)]
        s2 = np.where(reply, d[ref], s)
        d2 = np.where(reply, s[ref], d)
        src = s2 + 1
        dst = d2 + 1

    gaps = rng.exponential(shape.span_days * DAY / max(shape.n_edges, 1), size=E)
    ts = np.floor(np.cumsum(gaps)).astype(np.float64)
    return Stream(shape=shape,
                  src=src.astype(np.int32), dst=dst.astype(np.int32), ts=ts,
                  eidx=np.arange(1, E + 1, dtype=np.int64),
                  dst_pool=np.unique(dst).astype(np.int32))


def make_negatives(stream: Stream, batch_size: int, seed: int = 7) -> np.ndarray:
    """Per-batch negative destinations drawn like the reference's seeded sampler
    (/root/reference/utils/util.py:69-86: two randint draws per batch, second indexes
    the unique-destination pool)."""
    rs = np.random.RandomState(seed)
    E = stream.n_edges
    neg = np.empty(E, dtype=np.int32)
    n_src = len(np.unique(stream.src))
    pool = stream.dst_pool
    for lo in range(0, E, batch_size):
        size = min(batch_size, E - lo)
        rs.randint(0, n_src, size)
        neg[lo:lo + size] = pool[rs.randint(0, len(pool), size)]
    return neg


def edge_features(stream: Stream, seed: int = 11) -> np.ndarray:
    """[E+1, d_e] float32 with zero row 0 (d_e=0 -> the reference's zeros((E+1,1)))."""
    E, d = stream.n_edges, stream.shape.edge_dim
    if d == 0:
        return np.zeros((E + 1, 1), dtype=np.float32)
    rng = np.random.default_rng(seed)
    f = rng.standard_normal((E + 1, d), dtype=np.float32)
    f[0] = 0
    return f


def write_ml_format(stream: Stream, data_root: str, name: str | None = None, compact: bool = True) -> str:
    """Write `{data_root}/{name}/ml_{name}.csv|.npy` as the reference loader expects.

    The reference sizes its tables by the NUMBER of distinct node ids
    (/root/reference/utils/data_processing.py:97-99, train.py:124), i.e. it assumes ids 1..n without
    holes as its preprocessing produces; `compact` relabels the ids that occur, order-preserving."""
    name = name or stream.shape.name
    d = os.path.join(data_root, name)
    os.makedirs(d, exist_ok=True)
    E = stream.n_edges
    src, dst = stream.src, stream.dst
    if compact:
        ids, inv = np.unique(np.concatenate([src, dst]), return_inverse=True)
        src, dst = inv[:E] + 1, inv[E:] + 1
    with open(os.path.join(d, f"ml_{name}.csv"), "w") as fh:
        fh.write(",u,i,ts,label,idx\n")
        rows = np.column_stack([np.arange(E), src, dst])
        for r, t, e in zip(rows, stream.ts, stream.eidx):
            fh.write(f"{r[0]},{r[1]},{r[2]},{t},0,{e}\n")
    if stream.shape.edge_dim > 0:
        np.save(os.path.join(d, f"ml_{name}.npy"), edge_features(stream).astype(np.float64))
    return d


def wave_stats(src: np.ndarray, dst: np.ndarray, neg: np.ndarray | None = None) -> dict:
    """Depth of the read/write dependency DAG of one span (critical-path length in edges)."""
    last_w: dict[int, int] = {}
    last_r: dict[int, int] = {}
    depth = 0
    for i in range(len(src)):
        s, d = int(src[i]), int(dst[i])
        lvl = max(1 + last_w.get(s, -1), 1 + last_w.get(d, -1), last_r.get(s, 0), last_r.get(d, 0))
        if neg is not None:
            n = int(neg[i])
            lvl = max(lvl, 1 + last_w.get(n, -1))
            last_r[n] = max(last_r.get(n, 0), lvl)
        last_w[s] = lvl
        last_w[d] = lvl
        depth = max(depth, lvl + 1)
    return {"edges": len(src), "depth": depth, "mean_width": len(src) / max(depth, 1)}


def make_communities(name: str, n_comm: int, batch_size: int, n_edges: int | None = None):
    """Weak-scaling workload: `n_comm` independent streams of shape `name` (different seeds) merged
    into one chronological stream over n_comm x the nodes.  Community c's node v becomes
    1 + c*n_nodes + (v-1), so every community is spread over all ranks of a node-sharded run
    (owner = id & (world-1)) and a fraction (world-1)/world of its edges crosses shards.  The hub
    degrees stay those of one community while nodes and edges grow n_comm-fold.
    Returns (Stream, negatives); negatives are drawn per community like make_negatives."""
    shape = SHAPES[name]
    parts, negs = [], []
    for c in range(n_comm):
        st = make_stream(name, seed=SEEDS.get(name, 1000) + 100 * c, n_edges=n_edges)
        ng = make_negatives(st, batch_size, seed=7 + c)
        off = c * shape.n_nodes
        parts.append((st.src + off, st.dst + off, st.ts))
        negs.append(ng + off)
    if n_comm == 1:
        st = make_stream(name, n_edges=n_edges)
        return st, negs[0]
    src = np.concatenate([p[0] for p in parts])
    dst = np.concatenate([p[1] for p in parts])
    ts = np.concatenate([p[2] for p in parts])
    neg = np.concatenate(negs)
    order = np.argsort(ts, kind="stable")
    src, dst, ts, neg = src[order].astype(np.int32), dst[order].astype(np.int32), ts[order], neg[order].astype(np.int32)
    E = len(src)
    big = dataclasses.replace(shape, name=f"{name}x{n_comm}", n_users=shape.n_users * n_comm, n_items=shape.n_items * n_comm,
                              n_edges=E)
    return Stream(shape=big, src=src, dst=dst, ts=ts, eidx=np.arange(1, E + 1, dtype=np.int64),
                  dst_pool=np.unique(dst).astype(np.int32)), neg
