"""Level-1 drop-in for `GraphDiffusionEmbedding.compute_embedding_tppr_ensemble`
(LuckyLYM/Zebra modules/embedding_module.py:217-278): same signature and return value, but the
top-k tables, the query outputs and the gathered/concatenated neighbour tensor never leave the GPU.

    T-PPR query       tppr_finder.streaming_topk_raw / NeighborFinder.get_pruned_topk_device
    gather + concat + time encoding + weight normalisation   ONE fused kernel (csrc/gather_device.cuh)
    transform / transform_source (nn.Linear x2 each)          the module's own torch layers (north_star)

The function is written against the reference module's attributes (`fc1`, `fc2`, `fc1_source`,
`fc2_source`, `time_encoder`, `edge_features`, `tppr_finder`, `neighbor_finder`, `k`, `n_tppr`,
`alpha_list`, `beta_list`, `width`, `depth`, `tppr_strategy`), so checkpoints stay interchangeable;
`patch(module_cls)` installs it (see zebra_b200/launch.py).
"""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np
import torch

from ._lib import check, lib


# backward of the memory-row gathers: "sorted" (default) sums every node's gradient rows in a fixed order
# (reproducible, like the reference's index_put backward); "atomic" scatter-adds with red.global (faster on
# very large batches, not reproducible in the last bits)
REPRODUCIBLE_BACKWARD = os.environ.get("ZEBRA_GATHER_BACKWARD", "sorted").lower() != "atomic"


# weight normalisation of the neighbour sum: the reference's torch ops (default, bit-identical) or the fused
# kernel's output (ZEBRA_WEIGHT_NORM=fused: saves four small launches per instance, last-bit differences)
EXACT_WEIGHT_NORM = os.environ.get("ZEBRA_WEIGHT_NORM", "torch").lower() != "fused"
# ZEBRA_VALIDATE_IDS=1: every fused gather is followed by zebra_gather_check (one synchronisation per batch), so
# an out-of-range node / edge id raises at once like the reference's torch indexing; off by default, where
# such a row is skipped and the sticky flag is reported by the next zebra_gather_check
VALIDATE_IDS = os.environ.get("ZEBRA_VALIDATE_IDS", "0") == "1"


# lazy memory update on the device (level 2): the pending set is selected by a kernel,
# only its rows go through the memory updater, and the gather reads them in place of a cloned memory
LAZY_MEMORY = os.environ.get("ZEBRA_LAZY_MEMORY", "0") == "1"


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _sorted_sources(topk, src_nodes, n_tppr, rows, k, with_src):
    """Sources of the memory-row gathers (slots in buffer order, then the src rows) stably sorted by node."""
    plane = rows * k
    keys = [topk[j * plane * 16: j * plane * 16 + plane * 4].view(torch.int32) for j in range(n_tppr)]
    if with_src:
        keys.append(src_nodes.to(torch.int32))
    return torch.sort(torch.cat(keys), stable=True)


class FusedGather(torch.autograd.Function):
    """X[n_tppr, R*k, Dm+De+Dt], wn[n_tppr, R, k], src_rows[R, Dm] from the raw top-k buffer.
    Only `memory` is differentiable (edge features and the time encoder are frozen in the reference:
    model/tgn_model.py:41, model/time_encoding.py:18-23)."""

    @staticmethod
    def forward(ctx, memory, topk, src_nodes, edge_features, time_w, time_b, n_tppr, rows, k):
        L = lib()
        memory = memory.contiguous()
        N, Dm = memory.shape
        Ee, De = edge_features.shape
        Dt = time_w.numel()
        D = Dm + De + Dt
        dev = memory.device
        X = torch.empty((n_tppr, rows * k, D), dtype=torch.float32, device=dev)
        wn = torch.empty((n_tppr, rows, k), dtype=torch.float32, device=dev)
        src_rows = torch.empty((rows, Dm), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            check(L.zebra_gather_fwd(_p(topk), n_tppr, rows, k, _p(memory), N, Dm, _p(edge_features), Ee, De,
                                     _p(time_w), _p(time_b), Dt, _p(src_nodes), _p(X), _p(wn), _p(src_rows), st))
            if VALIDATE_IDS:
                check(L.zebra_gather_check(st))        # synchronises; raises on an out-of-range id, as the reference's indexing would
        ctx.save_for_backward(topk, src_nodes, wn)
        ctx.shape = (n_tppr, rows, k, N, Dm, D)
        ctx.mark_non_differentiable(wn)
        return X, wn, src_rows

    @staticmethod
    def backward(ctx, gX, _gwn, gsrc):
        topk, src_nodes, wn = ctx.saved_tensors
        n_tppr, rows, k, N, Dm, D = ctx.shape
        dev = topk.device
        gmem = torch.zeros((N, Dm), dtype=torch.float32, device=dev)
        gX = gX.contiguous() if gX is not None else torch.zeros((n_tppr, rows * k, D), dtype=torch.float32, device=dev)
        gsrc = gsrc.contiguous() if gsrc is not None else None
        with torch.cuda.device(dev):
            st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            if REPRODUCIBLE_BACKWARD:
                # sources sorted (stably) by destination node, summed in that order: run-to-run reproducible,
                # like the reference's sort-based index_put backward
                skeys, order = _sorted_sources(topk, src_nodes, n_tppr, rows, k, gsrc is not None)
                check(lib().zebra_gather_bwd_sorted(n_tppr, rows, k, _p(gX), _p(wn), _p(gsrc), _p(skeys), _p(order), _p(gmem),
                                                    N, Dm, D, st))
            else:
                check(lib().zebra_gather_bwd(_p(topk), n_tppr, rows, k, _p(gX), _p(wn), _p(gsrc),
                                             _p(src_nodes) if gsrc is not None else None, _p(gmem), N, Dm, D, st))
        return gmem, None, None, None, None, None, None, None, None


class FusedGatherLazy(torch.autograd.Function):
    """FusedGather over a lazily updated memory: rows of the pending set (bitmap / wordprefix from
    zebra_pending_select) are read from `upd_rows` [U, Dm] — the memory updater's output for exactly those
    nodes — every other row from `memory`.  Only `upd_rows` is differentiable; its gradient is the one
    autograd hands the updater in the reference (memory_updater.py:83: the gradient of the cloned memory
    at unique_node_ids), bit for bit, without a [num_nodes, Dm] tensor in either direction."""

    @staticmethod
    def forward(ctx, memory, upd_rows, bitmap, wordprefix, topk, src_nodes, edge_features, time_w, time_b, n_tppr, rows, k):
        L = lib()
        memory = memory.contiguous()
        upd_rows = upd_rows.contiguous()
        N, Dm = memory.shape
        Ee, De = edge_features.shape
        Dt = time_w.numel()
        D = Dm + De + Dt
        dev = memory.device
        X = torch.empty((n_tppr, rows * k, D), dtype=torch.float32, device=dev)
        wn = torch.empty((n_tppr, rows, k), dtype=torch.float32, device=dev)
        src_rows = torch.empty((rows, Dm), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            check(L.zebra_gather_fwd_lazy(_p(topk), n_tppr, rows, k, _p(memory), N, Dm, _p(edge_features), Ee, De,
                                          _p(time_w), _p(time_b), Dt, _p(src_nodes), _p(X), _p(wn), _p(src_rows),
                                          _p(upd_rows), _p(bitmap), _p(wordprefix), st))
            if VALIDATE_IDS:
                check(L.zebra_gather_check(st))
        ctx.save_for_backward(topk, src_nodes, wn, bitmap, wordprefix)
        ctx.shape = (n_tppr, rows, k, N, Dm, D, upd_rows.shape[0])
        ctx.mark_non_differentiable(wn)
        return X, wn, src_rows

    @staticmethod
    def backward(ctx, gX, _gwn, gsrc):
        topk, src_nodes, wn, bitmap, wordprefix = ctx.saved_tensors
        n_tppr, rows, k, N, Dm, D, U = ctx.shape
        dev = topk.device
        gupd = torch.zeros((U, Dm), dtype=torch.float32, device=dev)
        gX = gX.contiguous() if gX is not None else torch.zeros((n_tppr, rows * k, D), dtype=torch.float32, device=dev)
        gsrc = gsrc.contiguous() if gsrc is not None else None
        with torch.cuda.device(dev):
            st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            skeys, order = _sorted_sources(topk, src_nodes, n_tppr, rows, k, gsrc is not None)
            check(lib().zebra_gather_bwd_lazy(n_tppr, rows, k, _p(gX), _p(wn), _p(gsrc), _p(skeys), _p(order), _p(bitmap),
                                              _p(wordprefix), _p(gupd), N, Dm, D, st))
        return None, gupd, None, None, None, None, None, None, None, None, None, None


def pending_select(topk, n_tppr, rows, k, flags_dev, num_nodes):
    """(ids int64[U] ascending, bitmap, wordprefix) of the pending set: the distinct neighbour ids of `topk`
    whose byte in `flags_dev` (uint8[num_nodes], Memory.nodes) is set.  One 4-byte D2H for the count."""
    dev = topk.device
    n_words = (num_nodes + 31) // 32
    bitmap = torch.empty(n_words, dtype=torch.int32, device=dev)
    wordprefix = torch.empty(n_words, dtype=torch.int32, device=dev)
    ids = torch.empty(max(1, min(num_nodes, n_tppr * rows * k)), dtype=torch.int64, device=dev)
    count = torch.empty(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        check(lib().zebra_pending_select(_p(topk), n_tppr, rows, k, _p(flags_dev), num_nodes, _p(bitmap), _p(wordprefix),
                                         _p(ids), _p(count), st))
    return ids[:int(count.item())], bitmap, wordprefix


def _flags_to_device(self, mem, dev):
    """Memory.nodes (host numpy bool[num_nodes], modules/memory.py:22) through a pinned staging buffer."""
    n = mem.nodes.shape[0]
    pin = getattr(self, "_zebra_flags_pin", None)
    if pin is None or pin.numel() != n:
        pin = self._zebra_flags_pin = torch.empty(n, dtype=torch.uint8).pin_memory()
        self._zebra_flags_dev = torch.empty(n, dtype=torch.uint8, device=dev)
    pin.numpy()[:] = mem.nodes                  # the previous batch's copy has completed (count.item() synchronised)
    self._zebra_flags_dev.copy_(pin, non_blocking=True)
    return self._zebra_flags_dev


def _query_raw(self, source_nodes, timestamps, edge_idxs):
    """The T-PPR query of either strategy as one device buffer [n_tppr][4][R*k]."""
    R = len(source_nodes)
    if self.tppr_strategy == "streaming":
        buf, rows = self.tppr_finder.streaming_topk_raw(source_nodes, timestamps, edge_idxs)
        assert rows == R
        return buf
    if self.tppr_strategy == "pruning":
        parts = []
        for alpha, beta in zip(self.alpha_list, self.beta_list):
            p4 = self.neighbor_finder.get_pruned_topk_device(source_nodes, timestamps, self.width, self.depth, alpha, beta, self.k)
            parts.extend(x.reshape(-1).view(torch.uint8) for x in p4)
        return torch.cat(parts)
    raise ValueError(f"unknown tppr_strategy {self.tppr_strategy!r}")


def compute_embedding_tppr_ensemble(self, memory, source_nodes, timestamps, edge_idxs, memory_updater, train):
    source_nodes = np.array(source_nodes, dtype=np.int32)
    R, k, n_tppr = len(source_nodes), self.k, self.n_tppr
    dev = self.edge_features.device

    # the reference brackets its (synchronous, CPU) query with time.time() (embedding_module.py:220-225); the
    # query here is enqueued asynchronously, so it is bracketed with CUDA events on the stream instead and the
    # elapsed device time is added up when t_tppr is read (train.py:230, once per epoch) — see _TpprTimer
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    ev0.record()
    topk = _query_raw(self, source_nodes, timestamps, edge_idxs)
    ev1.record()
    _timer(self).pending.append((ev0, ev1))
    plane = R * k

    lazy = None
    if train and LAZY_MEMORY:
        # level 2: pending set selected on the device; only its rows go through the reference's updater
        # (memory_updater.py:83), nothing of size [num_nodes, Dm] is cloned or zero-filled
        mem = memory
        memory = mem.memory
        idx, bitmap, wordprefix = pending_select(topk, n_tppr, R, k, _flags_to_device(self, mem, dev), memory.shape[0])
        if idx.numel() > 0:
            upd = memory_updater.memory_updater(mem.messages[idx], memory[idx])
            lazy = (upd, bitmap, wordprefix)
    elif train:
        # the lazy memory update (modules/memory_updater.py:61-90) wants the distinct
        # neighbour ids on the host: one D2H of the id planes (n_tppr*R*k*4 bytes)
        ids = torch.stack([topk[j * plane * 16: j * plane * 16 + plane * 4].view(torch.int32) for j in range(n_tppr)])
        index = np.unique(ids.cpu().numpy())
        memory, _ = memory_updater.get_updated_memory(memory, index)

    n_edge = R // 3
    w0 = topk[plane * 12: plane * 16].view(torch.float32).view(R, k)
    self.average_topk = w0[:2 * n_edge].sum(dim=1).mean()        # stays on the device; nothing reads it per batch

    src_dev = torch.from_numpy(source_nodes).to(dev, non_blocking=True)
    tw = self.time_encoder.w.weight.detach().reshape(-1).contiguous()
    tb = self.time_encoder.w.bias.detach().contiguous()
    if lazy is not None:
        X, wn, src_rows = FusedGatherLazy.apply(memory, lazy[0], lazy[1], lazy[2], topk, src_dev, self.edge_features, tw, tb,
                                                n_tppr, R, k)
    else:
        X, wn, src_rows = FusedGather.apply(memory, topk, src_dev, self.edge_features, tw, tb, n_tppr, R, k)

    embeddings = self.transform_source(src_rows)
    for j in range(n_tppr):
        h = self.transform(X[j]).view(R, k, -1)
        if EXACT_WEIGHT_NORM:
            # the reference's own three ops (embedding_module.py:268-271) on the device-resident weights:
            # 12,000 floats per instance, and bit-identical to the reference whatever order torch.sum adds in
            # (the kernel's wn sums left to right and differs in the last bit for about half of the rows)
            wj = topk[j * plane * 16 + plane * 12: (j + 1) * plane * 16].view(torch.float32).view(R, k)
            ws = torch.sum(wj, dim=1)
            wnj = wj / ws.unsqueeze(1)
            wnj[ws == 0] = 0
        else:
            wnj = wn[j]
        h = torch.sum(h * wnj[:, :, None], dim=1)
        embeddings = torch.cat((embeddings, h), dim=1)
    return embeddings


class _TpprTimer:
    """Seconds spent in the T-PPR query: CUDA-event time of every query since the last reset, summed when
    read.  An event pair spans from the moment the stream reaches the query (or the host starts enqueuing it,
    whichever is later) to the end of its last kernel — host dependency pass, H2D and kernel included."""

    def __init__(self):
        self.base, self.pending = 0.0, []

    def read(self):
        if self.pending:
            self.pending[-1][1].synchronize()
            self.base += sum(a.elapsed_time(b) for a, b in self.pending) * 1e-3
            self.pending = []
        return self.base


def _timer(self):
    t = self.__dict__.get("_zebra_tppr_timer")
    if t is None:
        t = self.__dict__["_zebra_tppr_timer"] = _TpprTimer()
    return t


def _get_t_tppr(self):
    return _timer(self).read()


def _set_t_tppr(self, value):
    t = _timer(self)
    t.base, t.pending = float(value), []


def patch(module_cls):
    """Install the device-resident path on the reference's GraphDiffusionEmbedding class."""
    module_cls.compute_embedding_tppr_ensemble = compute_embedding_tppr_ensemble
    module_cls.t_tppr = property(_get_t_tppr, _set_t_tppr)      # read by train.py:230, reset by model/tgn_model.py:121
    return module_cls


class TimeEncode(torch.nn.Module):
    """Same parameters as the reference's TimeEncode (model/time_encoding.py:6-28): w_j = 10^(-9j/(d-1)),
    zero bias, both frozen.  The forward pass itself runs inside the fused gather kernel."""

    def __init__(self, dimension):
        super().__init__()
        self.dimension = dimension
        self.w = torch.nn.Linear(1, dimension)
        self.w.weight = torch.nn.Parameter(torch.from_numpy(1 / 10 ** np.linspace(0, 9, dimension, dtype=np.float32)).reshape(dimension, -1))
        self.w.bias = torch.nn.Parameter(torch.zeros(dimension))
        self.w.weight.requires_grad = False
        self.w.bias.requires_grad = False

    @torch.no_grad()
    def forward(self, t):
        return torch.cos(self.w(t.unsqueeze(dim=2)))


class GraphDiffusionEmbedding(torch.nn.Module):
    """Host-side mirror of the reference's GraphDiffusionEmbedding (modules/embedding_module.py:76-136,
    217-331) for users who do not run the reference's train.py: same parameter names (its state_dict
    loads here and vice versa), same T-PPR protocol methods, device-resident query + fused gather."""

    t_tppr = property(_get_t_tppr, _set_t_tppr)       # CUDA-event time of the queries (see _TpprTimer)

    def __init__(self, edge_features, neighbor_finder, n_edge_features, n_time_features, embedding_dimension, device,
                 args, num_nodes):
        super().__init__()
        self.edge_features = edge_features
        self.neighbor_finder = neighbor_finder
        self.time_encoder = TimeEncode(n_time_features)
        self.device = device
        self.args = args
        self.num_nodes = num_nodes
        self.t_tppr = 0
        d = embedding_dimension
        self.fc1 = torch.nn.Linear(d + n_time_features + n_edge_features, d)
        self.fc2 = torch.nn.Linear(d, d)
        self.act = torch.nn.ReLU()
        self.drop = torch.nn.Dropout(0.1)
        self.fc1_source = torch.nn.Linear(d, d)
        self.fc2_source = torch.nn.Linear(d, d)
        self.combiner = torch.nn.Linear(d + d, d)
        for lin in (self.fc1, self.fc2, self.fc1_source, self.fc2_source):
            torch.nn.init.xavier_normal_(lin.weight)
        self.n_tppr = len(args.alpha_list)
        self.alpha_list, self.beta_list = args.alpha_list, args.beta_list
        self.k = args.topk
        self.tppr_strategy = args.tppr_strategy
        self.width, self.depth = args.n_degree, args.n_layer
        assert self.k != 0
        if self.tppr_strategy == "streaming":
            from .finder import tppr_finder
            self.tppr_finder = tppr_finder(num_nodes, self.k, self.n_tppr, self.alpha_list, self.beta_list, device=device)

    def reset_tppr(self):
        self.tppr_finder.reset_tppr()

    def backup_tppr(self):
        return self.tppr_finder.backup_tppr()

    def restore_tppr(self, backup):
        self.tppr_finder.restore_tppr(backup)

    def streaming_topk(self, source_nodes, timestamps, edge_idxs):
        return self.tppr_finder.streaming_topk(source_nodes, timestamps, edge_idxs)

    def streaming_topk_no_fake(self, source_nodes, timestamps, edge_idxs):
        return self.tppr_finder.streaming_topk_no_fake(source_nodes, timestamps, edge_idxs)

    def fill_tppr(self, sources, targets, timestamps, edge_idxs, tppr_filled):
        if tppr_filled:
            self.tppr_finder.restore_val_tppr()
        else:
            self.tppr_finder.compute_val_tppr(sources, targets, timestamps, edge_idxs)

    def pruning_topk(self, source_nodes, timestamps):
        n = len(source_nodes)
        outs = [[np.zeros((n, self.k), dt) for _ in range(self.n_tppr)] for dt in (np.int32, np.int32, np.float32, np.float32)]
        for i, (alpha, beta) in enumerate(zip(self.alpha_list, self.beta_list)):
            self.neighbor_finder.get_pruned_topk(source_nodes, timestamps, self.width, self.depth, alpha, beta, self.k,
                                                 outs[0][i], outs[1][i], outs[2][i], outs[3][i])
        return tuple(outs)

    def transform_source(self, x):
        return self.fc2_source(self.drop(self.act(self.fc1_source(x))))

    def transform(self, x):
        return self.fc2(self.drop(self.act(self.fc1(x))))

    def combine(self, x):
        return self.combiner(x)

    compute_embedding_tppr_ensemble = compute_embedding_tppr_ensemble
