"""CPU oracle for the T-PPR hot path — TEST INFRASTRUCTURE, NOT PRODUCT.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  zebra_b200/ never does.

`OracleFinder` mirrors the reference's `tppr_finder` interface (/root/reference/utils/util.py:
391-873) on top of oracle/tppr_oracle.c, including the reference's de-facto aliasing behaviour of
backup_tppr / restore_tppr / restore_val_tppr (util.py:436-444, 872-873: numba typed-List
`.copy()` is shallow, so the "backups" alias the live containers — SURVEY.md H1).
`OracleNeighborFinder` mirrors `NeighborFinder.get_pruned_topk` (util.py:185-276).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libtppr_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "tppr_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", _SO, src, "-lm"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.zo_table_new.restype = C.c_void_p
        L.zo_table_new.argtypes = [C.c_int64, C.c_int32]
        L.zo_table_free.argtypes = [C.c_void_p]
        L.zo_table_clone.restype = C.c_void_p
        L.zo_table_clone.argtypes = [C.c_void_p]
        L.zo_table_get.restype = C.c_int32
        L.zo_table_get.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 5
        L.zo_stats_get.argtypes = [C.c_void_p]
        L.zo_stream.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_int64, C.c_int32, C.c_int32] + [C.c_void_p] * 4
        L.zo_pruned_topk.argtypes = [C.c_void_p] * 6 + [C.c_int64, C.c_int32, C.c_int32, C.c_double,
                                                         C.c_double, C.c_int32, C.c_int32] + [C.c_void_p] * 4
        L.zo_numba_argsort.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def numba_argsort(values: np.ndarray) -> np.ndarray:
    v = np.ascontiguousarray(values, dtype=np.float64)
    r = np.empty(len(v), dtype=np.int32)
    lib().zo_numba_argsort(_p(v), len(v), _p(r))
    return r


class _Table:
    def __init__(self, num_nodes, k, ptr=None):
        self.num_nodes, self.k = num_nodes, k
        self.ptr = ptr if ptr is not None else lib().zo_table_new(num_nodes, k)

    def clone(self):
        return _Table(self.num_nodes, self.k, lib().zo_table_clone(self.ptr))

    def get(self, v):
        k = self.k
        e = np.empty(k, np.int64); n = np.empty(k, np.int64)
        t = np.empty(k, np.float64); s = np.empty(k, np.float64)
        nm = np.zeros(1, np.float64)
        c = lib().zo_table_get(self.ptr, int(v), _p(e), _p(n), _p(t), _p(s), _p(nm))
        return e[:c].copy(), n[:c].copy(), t[:c].copy(), s[:c].copy(), float(nm[0])

    def __del__(self):
        try:
            lib().zo_table_free(self.ptr)
        except Exception:
            pass


class OracleFinder:
    """Same surface as the reference's tppr_finder; `tie_mode` 0 = numba argsort, 1 = stable."""

    def __init__(self, num_nodes, k, n_tppr, alpha_list, beta_list, tie_mode=0, strict_snapshots=False):
        self.num_nodes, self.k, self.n_tppr = int(num_nodes), int(k), int(n_tppr)
        self.alpha_list, self.beta_list = list(alpha_list), list(beta_list)
        self.tie_mode = tie_mode
        self.strict = strict_snapshots
        self.reset_val_tppr()
        self.reset_tppr()

    def reset_tppr(self):
        self.tables = [_Table(self.num_nodes, self.k) for _ in range(self.n_tppr)]

    def reset_val_tppr(self):
        self.val_tables = [_Table(self.num_nodes, self.k) for _ in range(self.n_tppr)]

    def _snap(self, tables):
        return [t.clone() for t in tables] if self.strict else list(tables)

    def backup_tppr(self):
        return self._snap(self.tables)

    def restore_tppr(self, backup):
        self.tables = list(backup)

    def restore_val_tppr(self):
        self.tables = self._snap(self.val_tables)

    def _run(self, nodes, ts, eidx, rows_per_edge, instances):
        n = len(eidx)
        nodes = np.ascontiguousarray(nodes, dtype=np.int64)
        ts = np.ascontiguousarray(ts, dtype=np.float64)
        eidx = np.ascontiguousarray(eidx, dtype=np.int64)
        outs = ([], [], [], [])
        for j in range(self.n_tppr):
            rows = rows_per_edge * n
            o = (np.zeros((rows, self.k), np.int32), np.zeros((rows, self.k), np.int32),
                 np.zeros((rows, self.k), np.float32), np.zeros((rows, self.k), np.float32))
            if j in instances:
                lib().zo_stream(self.tables[j].ptr, self.alpha_list[j], self.beta_list[j], _p(nodes), _p(ts),
                                _p(eidx), n, rows_per_edge, self.tie_mode, *[_p(a) for a in o])
            for lst, a in zip(outs, o):
                lst.append(a)
        return outs

    def streaming_topk(self, source_nodes, timestamps, edge_idxs):
        return self._run(source_nodes, timestamps, edge_idxs, 3, range(self.n_tppr))

    def streaming_topk_no_fake(self, source_nodes, timestamps, edge_idxs):
        return self._run(source_nodes, timestamps, edge_idxs, 2, range(self.n_tppr))

    def single_streaming_topk(self, source_nodes, timestamps, edge_idxs, tppr_id):
        o = self._run(source_nodes, timestamps, edge_idxs, 3, [tppr_id])
        return tuple(x[tppr_id] for x in o)

    def compute_val_tppr(self, sources, targets, timestamps, edge_idxs):
        nodes = np.concatenate([np.asarray(sources, np.int64), np.asarray(targets, np.int64)])
        self._run(nodes, timestamps, edge_idxs, 0, range(self.n_tppr))
        self.val_tables = self._snap(self.tables)

    def state(self, inst, v):
        return self.tables[inst].get(v)


def build_csr(src, dst, eidx, ts, num_nodes):
    """Time-sorted adjacency as CSR, the oracle-side equivalent of get_neighbor_finder
    (/root/reference/utils/util.py:90-107: both directions, stable sort by timestamp)."""
    src = np.asarray(src, np.int64); dst = np.asarray(dst, np.int64)
    owner = np.concatenate([src, dst]); other = np.concatenate([dst, src])
    # python's sorted() is stable over the insertion order, which interleaves (src,dst) per edge
    order_in = np.concatenate([2 * np.arange(len(src)), 2 * np.arange(len(src)) + 1])
    e2 = np.concatenate([eidx, eidx]); t2 = np.concatenate([ts, ts]).astype(np.float64)
    perm = np.lexsort((order_in, t2, owner))
    indptr = np.zeros(num_nodes + 1, np.int64)
    np.add.at(indptr, owner + 1, 1)
    indptr = np.cumsum(indptr)
    return (indptr, other[perm].astype(np.int32), np.asarray(e2)[perm].astype(np.int32), t2[perm])


class OracleNeighborFinder:
    def __init__(self, indptr, adj_node, adj_eidx, adj_ts, tie_mode=0):
        self.indptr = np.ascontiguousarray(indptr, np.int64)
        self.adj_node = np.ascontiguousarray(adj_node, np.int32)
        self.adj_eidx = np.ascontiguousarray(adj_eidx, np.int32)
        self.adj_ts = np.ascontiguousarray(adj_ts, np.float64)
        self.tie_mode = tie_mode

    def get_pruned_topk(self, source_nodes, timestamps, width, depth, alpha, beta, k,
                        node_list, edge_idxs_list, delta_time_list, weight_list):
        q = np.ascontiguousarray(source_nodes, np.int32)
        t = np.ascontiguousarray(timestamps, np.float64)
        for a in (node_list, edge_idxs_list, delta_time_list, weight_list):
            assert a.flags.c_contiguous
        lib().zo_pruned_topk(_p(self.indptr), _p(self.adj_node), _p(self.adj_eidx), _p(self.adj_ts), _p(q), _p(t),
                             len(q), width, depth, alpha, beta, k, self.tie_mode,
                             _p(node_list), _p(edge_idxs_list), _p(delta_time_list), _p(weight_list))


def stats():
    o = np.zeros(4, np.int64)
    lib().zo_stats_get(_p(o))
    return {"prunes": int(o[0]), "prunes_with_tie": int(o[1]), "boundary_ties": int(o[2]), "relevant_ties": int(o[3])}


def stats_reset():
    lib().zo_stats_reset()
