"""Drop-in replacement for the reference's `NeighborFinder` / `get_neighbor_finder`
(LuckyLYM/Zebra utils/util.py:90-107, 144-276): the static time-sorted adjacency lives in HBM as CSR
and `get_pruned_topk` (the `--tppr_strategy pruning` query, util.py:185-276) runs in the sm_100a
kernel `prune_topk_kernel` behind the C ABI (include/zebra_tppr.h).  Same constructor arguments,
method names, argument meaning and in-place output convention as the reference.  No CPU path for
the query: without a CUDA device construction raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ._lib import check, lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def build_csr(sources, destinations, edge_idxs, timestamps, num_nodes=None):
    """Time-sorted adjacency of an undirected temporal edge list as CSR, in the order the reference
    builds it (util.py:90-107): every edge is appended to both endpoints' lists in stream order and
    each list is then stably sorted by timestamp."""
    src = np.asarray(sources, np.int64)
    dst = np.asarray(destinations, np.int64)
    E = len(src)
    if num_nodes is None:
        num_nodes = int(max(src.max(), dst.max())) + 1 if E else 1
    owner = np.concatenate([src, dst])
    other = np.concatenate([dst, src])
    seq = np.concatenate([2 * np.arange(E), 2 * np.arange(E) + 1])      # append order within a list
    e2 = np.concatenate([np.asarray(edge_idxs)] * 2)
    t2 = np.concatenate([np.asarray(timestamps, np.float64)] * 2)
    perm = np.lexsort((seq, t2, owner))
    indptr = np.zeros(num_nodes + 1, np.int64)
    np.add.at(indptr, owner + 1, 1)
    np.cumsum(indptr, out=indptr)
    return (indptr, np.ascontiguousarray(other[perm], np.int32), np.ascontiguousarray(e2[perm], np.int32),
            np.ascontiguousarray(t2[perm], np.float64))


class NeighborFinder:  # noqa: N801 (name kept from the reference)
    def __init__(self, node_to_neighbors, node_to_edge_idxs, node_to_edge_timestamps, device=None, _csr=None):
        if not torch.cuda.is_available():
            raise RuntimeError("zebra_b200.NeighborFinder needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        if _csr is None:
            lens = np.fromiter((len(a) for a in node_to_neighbors), np.int64, len(node_to_neighbors))
            indptr = np.zeros(len(lens) + 1, np.int64)
            np.cumsum(lens, out=indptr[1:])
            cat = lambda lst, dt: (np.concatenate([np.asarray(a, dt) for a in lst]) if indptr[-1] else np.zeros(0, dt))  # noqa: E731
            _csr = (indptr, cat(node_to_neighbors, np.int32), cat(node_to_edge_idxs, np.int32),
                    cat(node_to_edge_timestamps, np.float64))
        self.indptr, self.adj_node, self.adj_eidx, self.adj_ts = (np.ascontiguousarray(a) for a in _csr)
        self.num_nodes = len(self.indptr) - 1
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._L = lib()
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(self._L.zebra_csr_create(C.byref(h), _ptr(self.indptr), _ptr(self.adj_node), _ptr(self.adj_eidx),
                                           _ptr(self.adj_ts), self.num_nodes, len(self.adj_node), self.device.index or 0))
        self._h = h
        self._host_out = None

    @classmethod
    def from_edges(cls, sources, destinations, edge_idxs, timestamps, num_nodes=None, device=None):
        return cls(None, None, None, device=device, _csr=build_csr(sources, destinations, edge_idxs, timestamps, num_nodes))

    # ---- host-side views, as in the reference (util.py:152-173) ----
    def find_before(self, src_idx, cut_time):
        lo, hi = self.indptr[src_idx], self.indptr[src_idx + 1]
        i = int(np.searchsorted(self.adj_ts[lo:hi], cut_time))
        return self.adj_node[lo:lo + i], self.adj_eidx[lo:lo + i], self.adj_ts[lo:lo + i]

    # ---- the pruning-strategy query (util.py:185-276) ----
    def _query(self, source_nodes, timestamps, width, depth, alpha, beta, k, to_host=True):
        q = np.ascontiguousarray(source_nodes, dtype=np.int32)
        t = np.ascontiguousarray(timestamps, dtype=np.float64)
        nq = len(q)
        if len(t) < nq:
            raise ValueError("timestamps shorter than source_nodes")
        nbytes = self._L.zebra_prune_out_bytes(nq, int(k))
        dev_out = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=self.device)
        host = None
        if to_host:
            if self._host_out is None or self._host_out.numel() < nbytes:
                self._host_out = torch.empty(max(nbytes, 1), dtype=torch.uint8).pin_memory()
            host = self._host_out
        with torch.cuda.device(self.device):
            st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            check(self._L.zebra_prune_topk(self._h, _ptr(q), _ptr(t), nq, int(width), int(depth), float(alpha),
                                           float(beta), int(k), C.c_void_p(dev_out.data_ptr()),
                                           C.c_void_p(host.data_ptr()) if to_host else None, st))
        plane = nq * k
        if to_host:
            buf = host.numpy()[:nbytes]
            parts = [buf[i * plane * 4:(i + 1) * plane * 4].view(dt).reshape(nq, k)
                     for i, dt in enumerate((np.int32, np.int32, np.float32, np.float32))]
            filled = buf[4 * plane * 4:4 * plane * 4 + nq * 4].view(np.int32).astype(bool)
            return parts, filled
        parts = [dev_out[i * plane * 4:(i + 1) * plane * 4].view(dt).view(nq, k)
                 for i, dt in enumerate((torch.int32, torch.int32, torch.float32, torch.float32))]
        filled = dev_out[4 * plane * 4:4 * plane * 4 + nq * 4].view(torch.int32)
        return parts, filled

    def get_pruned_topk(self, source_nodes, timestamps, width, depth, alpha, beta, k,
                        node_list, edge_idxs_list, delta_time_list, weight_list):
        """Writes the top-k rows into the four caller arrays [len(source_nodes), k] in place; rows of
        queries without any earlier neighbour are left untouched, as the reference does."""
        parts, filled = self._query(source_nodes, timestamps, width, depth, alpha, beta, k)
        for dst, src in zip((node_list, edge_idxs_list, delta_time_list, weight_list), parts):
            if filled.all():
                dst[...] = src
            else:
                dst[filled] = src[filled]

    def get_pruned_topk_device(self, source_nodes, timestamps, width, depth, alpha, beta, k):
        """Level-1 fast path: (nodes, eidx, dt, w) as CUDA tensors [n, k] (rows of empty queries are
        zero, which is what the reference's pre-zeroed arrays hold), asynchronous on the current stream."""
        parts, _ = self._query(source_nodes, timestamps, width, depth, alpha, beta, k, to_host=False)
        return parts

    def launches(self):
        return int(self._L.zebra_prune_launches(self._h))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._L.zebra_csr_destroy(self._h)
                self._h = None
        except Exception:
            pass


def get_neighbor_finder(data, device=None):
    """Same role as the reference's factory (util.py:90-107): `data` has sources, destinations,
    edge_idxs, timestamps."""
    return NeighborFinder.from_edges(data.sources, data.destinations, data.edge_idxs, data.timestamps, device=device)
