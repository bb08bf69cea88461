"""Node-sharded T-PPR across the GPUs of one NVSwitch box (include/zebra_tppr.h "node-sharded mode").

One process per GPU (torchrun) — or, for tests, several handles inside one process.  T-PPR state is
partitioned by node id (node v lives on rank v & (world-1)); every rank streams the same replicated
edge span through the same plan; the span kernel exchanges cross-shard endpoint vectors through
peer-mapped mailboxes (NVLink stores issued by the kernel itself) and every owner writes its rows
into the output region of every rank.  The result on each rank is byte-identical to the single-GPU
`tppr_finder` (tests/test_gpu_sharded.py, tests/test_gpu_sharded_mp.py).

Host-side plumbing only lives here: `torch.distributed` carries the 64-byte CUDA IPC handles and the
barriers; there is no data-path collective.  The pure-host helpers (`owner_of`, `exchange_stats`,
`gather_handles`) are covered by world_size-2 gloo tests on CPU (tests/test_sharded_host.py).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

HANDLE_BYTES = 64


def owner_of(nodes, world: int):
    """Rank that holds each node's vector: v & (world-1); world must be a power of two."""
    if world < 1 or world & (world - 1):
        raise ValueError("world must be a power of two")
    return np.asarray(nodes) & (world - 1)


def local_row(nodes, world: int):
    return np.asarray(nodes) >> (int(world).bit_length() - 1)


def exchange_stats(src, dst, neg, world: int, k: int, n_tppr: int, row_mode: int = 0):
    """What the sharded kernel moves over NVLink for a span, computed from the edge list alone:
    cross-shard edges push two old vectors per instance (24k+8 algorithmic bytes each, SURVEY.md
    §8d), and every output row (16k bytes per instance) is stored to the world-1 other ranks
    (row_mode 0) or to its one consumer rank, remote with probability (world-1)/world (row_mode 1)."""
    src, dst = np.asarray(src), np.asarray(dst)
    os_, od = owner_of(src, world), owner_of(dst, world)
    cross = int(np.count_nonzero((os_ != od) & (src != dst)))
    rows = len(src) * (3 if neg is not None else 2)
    per_rank_updates = np.bincount(np.concatenate([os_, od[src != dst]]), minlength=world)
    return {
        "edges": int(len(src)), "cross_edges": cross, "cross_fraction": cross / max(1, len(src)),
        "mailbox_bytes": cross * 2 * n_tppr * (24 * k + 8),
        "row_bytes": rows * n_tppr * 16 * k * (world - 1) if row_mode == 0 else rows * n_tppr * 16 * k * (world - 1) // world,
        "updates_per_rank": per_rank_updates.tolist(),
    }


def gather_handles(my_handle: bytes, group=None):
    """All ranks' 64-byte IPC handles, in rank order, through torch.distributed (any backend)."""
    import torch
    import torch.distributed as dist
    if len(my_handle) != HANDLE_BYTES:
        raise ValueError("an IPC handle is 64 bytes")
    world = dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    mine = torch.tensor(list(my_handle), dtype=torch.uint8, device=dev)
    out = [torch.empty(HANDLE_BYTES, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return [bytes(t.cpu().tolist()) for t in out]


class _DevBuf:
    """Library-owned device memory exposed to torch through __cuda_array_interface__."""

    def __init__(self, ptr, nbytes, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 3}


class ShardedFinder:
    """One rank's share of a node-sharded `tppr_finder`.  Same constructor arguments as the
    reference's tppr_finder (utils/util.py:393-400) plus the shard geometry."""

    def __init__(self, num_nodes, k, n_tppr, alpha_list, beta_list, rank, world, max_window_edges, out_bytes,
                 device=None):
        import torch
        from ._lib import check, lib
        if not torch.cuda.is_available():
            raise RuntimeError("zebra_b200.ShardedFinder needs CUDA devices (B200, sm_100a); there is no CPU fallback")
        self.num_nodes, self.k, self.n_tppr = int(num_nodes), int(k), int(n_tppr)
        self.alpha_list = [float(a) for a in alpha_list]
        self.beta_list = [float(b) for b in beta_list]
        self.rank, self.world = int(rank), int(world)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._L, self._check = lib(), check
        h = C.c_void_p()
        a = (C.c_double * self.n_tppr)(*self.alpha_list)
        b = (C.c_double * self.n_tppr)(*self.beta_list)
        check(self._L.zebra_tppr_create(C.byref(h), self.num_nodes, self.k, self.n_tppr, a, b, self.device.index or 0))
        self._h = h
        buf = (C.c_uint8 * HANDLE_BYTES)()
        check(self._L.zebra_tppr_shard_init(h, self.rank, self.world, int(max_window_edges), int(out_bytes), buf))
        self.ipc_handle = bytes(buf)
        self.max_window_edges = int(max_window_edges)
        self._record_bytes = self._L.zebra_tppr_record_bytes(h)
        nbytes = int(self._L.zebra_tppr_shard_out_bytes(h))
        with torch.cuda.device(self.device):
            self.out = torch.as_tensor(_DevBuf(self._L.zebra_tppr_shard_out(h), nbytes, self), device=self.device)

    # ---- wiring ----
    def connect(self, handles):
        """Multi-process: `handles` = every rank's ipc_handle in rank order (gather_handles)."""
        blob = b"".join(handles)
        if len(blob) != HANDLE_BYTES * self.world:
            raise ValueError("need one 64-byte handle per rank")
        self._check(self._L.zebra_tppr_shard_connect(self._h, C.c_char_p(blob)))

    @staticmethod
    def connect_local(finders):
        """Same process: all ranks' finders (possibly on one GPU, see set_max_blocks)."""
        arr = (C.c_void_p * len(finders))(*[f._h for f in finders])
        for f in finders:
            f._check(f._L.zebra_tppr_shard_connect_local(f._h, arr))

    def set_row_mode(self, mode):
        """0: every rank receives every row (default); 1: the rows of batch b go to rank b & (world-1)."""
        self._check(self._L.zebra_tppr_shard_row_mode(self._h, int(mode)))
        self.row_mode = int(mode)

    def set_max_blocks(self, blocks):
        self._check(self._L.zebra_tppr_set_max_blocks(self._h, int(blocks)))

    # ---- reference surface that makes sense per shard ----
    def _stream_ptr(self):
        import torch
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def reset_tppr(self):
        """Zero this rank's table slices (utils/util.py:402-417).  The caller barriers across ranks
        before the next launch."""
        import torch
        with torch.cuda.device(self.device):
            self._check(self._L.zebra_tppr_shard_reset(self._h, self._stream_ptr()))

    def make_plan(self, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge=3):
        from .finder import StreamPlan
        return _ShardedPlan(self, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge)

    def epoch(self, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge=3, window=None, windows=False):
        """Pipelined pass over a chronological span held in HOST memory (include/zebra_tppr.h "pipelined
        epochs" on a node-sharded handle): every rank calls it with the same replicated arrays.  Windows of
        `window` batches (a multiple of the world size) are staged (ownership-aware dependency pass on worker
        threads), computed by the sharded span kernel and — once every rank has finished the window — this
        rank's batches (b % world == rank under row_mode 1, all batches under row_mode 0) are copied to pinned
        host memory.  Returns an iterator that yields (batch index, nodes, eidx, dt, w), each a list of n_tppr
        numpy arrays [rows, k] exactly like `tppr_finder.streaming_topk` of that batch (views of pinned library
        memory, valid until the iterator moves past their window).  `windows=True` yields one
        (batch indices, nodes, eidx, dt, w) tuple per window instead, each array [n_own_batches, rows, k] (full
        batches; a ragged last batch of the span comes as its own tuple)."""
        return _ShardedEpoch(self, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge, window, windows)

    def state(self, inst, nodes):
        """Vectors of `nodes` (all owned by this rank), like tppr_finder.state."""
        import torch
        from .finder import _ptr
        ids = np.ascontiguousarray(nodes, dtype=np.int64)
        m, k = len(ids), self.k
        cnt = np.zeros(m, np.int32); e = np.zeros((m, k), np.int32); nd = np.zeros((m, k), np.int32)
        t = np.zeros((m, k), np.float64); s = np.zeros((m, k), np.float64); nm = np.zeros(m, np.float64)
        with torch.cuda.device(self.device):
            self._check(self._L.zebra_tppr_read_state(self._h, int(inst), _ptr(ids), m, _ptr(cnt), _ptr(e), _ptr(nd),
                                                      _ptr(t), _ptr(s), _ptr(nm), self._stream_ptr()))
        return [(e[a, :cnt[a]].astype(np.int64), nd[a, :cnt[a]].astype(np.int64), t[a, :cnt[a]], s[a, :cnt[a]], float(nm[a]))
                for a in range(m)]

    def sync_check(self):
        self._check(self._L.zebra_tppr_sync_check(self._h, self._stream_ptr()))

    def counters(self):
        from .finder import _ptr
        o = np.zeros(8, np.int64)
        self._check(self._L.zebra_tppr_counters(self._h, _ptr(o)))
        return {"items": int(o[0]), "tie_prunes": int(o[1]), "launches": int(o[2]), "timeouts": int(o[3]),
                "merge_ranked": int(o[4]), "chain_items": int(o[5]), "chain_steps": int(o[6])}

    def last_kernel_ms(self):
        return float(self._L.zebra_tppr_last_kernel_ms(self._h))

    def disconnect(self):
        """Unmap the peers' arenas.  Multi-process teardown: every rank disconnects, barrier, close."""
        if getattr(self, "_h", None):
            self._check(self._L.zebra_tppr_shard_disconnect(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self.out = None
            self._L.zebra_tppr_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _make_sharded_plan_class():
    from .finder import StreamPlan

    class ShardedPlan(StreamPlan):
        """StreamPlan whose rows land in the shared output region of every rank (finder.out)."""

        def run(self, first=0, n_batches=None, batches_per_launch=1, out=None, out_slots=None, out_offset=0):
            """Rows go to slot (b - first) % out_slots of finder.out[out_offset:] on EVERY rank."""
            import torch
            if out is not None:
                raise ValueError("sharded plans write to finder.out")
            n_batches = self.n_batches - first if n_batches is None else n_batches
            if batches_per_launch * self.batch_size > self.f.max_window_edges:
                raise ValueError("launch window exceeds the mailbox window given to ShardedFinder")
            if out_slots is None:
                out_slots = min(batches_per_launch, n_batches)
            f = self.f
            with torch.cuda.device(f.device):
                f._check(f._L.zebra_tppr_plan_run(f._h, self._p, first, n_batches, batches_per_launch,
                                                  C.c_void_p(int(out_offset)) if out_offset else None, out_slots, f._stream_ptr()))
            return f.out[int(out_offset):] if out_offset else f.out

    return ShardedPlan


def _ShardedPlan(*a, **kw):
    return _make_sharded_plan_class()(*a, **kw)


def create_sharded(num_nodes, k, n_tppr, alpha_list, beta_list, max_window_edges, out_bytes, group=None, device=None,
                   row_mode=0):
    """torchrun entry: build this rank's ShardedFinder, exchange IPC handles, connect, barrier."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    f = ShardedFinder(num_nodes, k, n_tppr, alpha_list, beta_list, rank, world, max_window_edges, out_bytes, device=device)
    f.set_row_mode(row_mode)
    f.connect(gather_handles(f.ipc_handle, group))
    dist.barrier(group)
    return f


class _ShardedEpoch:
    def __init__(self, f, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge, window, windows=False):
        import torch
        from .finder import _ptr
        self.f = f
        self.src = np.ascontiguousarray(src, dtype=np.int32)
        self.dst = np.ascontiguousarray(dst, dtype=np.int32)
        self.neg = None if neg is None else np.ascontiguousarray(neg, dtype=np.int32)
        self.ts = np.ascontiguousarray(timestamps, dtype=np.float64)
        self.eidx = np.ascontiguousarray(edge_idxs, dtype=np.int64)
        self.n, self.bs, self.rows_per_edge = len(self.src), int(batch_size), int(rows_per_edge)
        if window is None:
            window = max(f.world, (f.max_window_edges // self.bs) // f.world * f.world)
        self.window = int(window)
        self.by_window = bool(windows)
        self._ep = C.c_void_p()
        with torch.cuda.device(f.device):
            f._check(f._L.zebra_tppr_epoch_begin(f._h, _ptr(self.src), _ptr(self.dst), _ptr(self.neg), _ptr(self.ts), _ptr(self.eidx),
                                                 self.n, self.rows_per_edge, self.bs, self.window, f._stream_ptr(), C.byref(self._ep)))
        self.slot = int(f._L.zebra_tppr_epoch_slot_bytes(self._ep))

    def __iter__(self):
        f, L = self.f, self.f._L
        rows_ptr, b0, nb = C.c_void_p(), C.c_int64(), C.c_int64()
        T, k, R = f.n_tppr, f.k, self.rows_per_edge * self.bs
        mode1 = getattr(f, "row_mode", 0) == 1
        try:
            while True:
                f._check(L.zebra_tppr_epoch_next(self._ep, C.byref(rows_ptr), C.byref(b0), C.byref(nb)))
                if nb.value == 0:
                    break
                if not self.rows_per_edge:
                    continue
                first, cnt = b0.value, nb.value
                mine = list(range(first + f.rank, first + cnt, f.world)) if mode1 else list(range(first, first + cnt))
                if not mine:
                    continue
                raw = np.ctypeslib.as_array(C.cast(rows_ptr, C.POINTER(C.c_uint8)), shape=(len(mine) * self.slot,))
                if self.by_window:
                    ragged = (self.n - mine[-1] * self.bs) < self.bs
                    full = len(mine) - (1 if ragged else 0)
                    if full:
                        wi = raw[:full * self.slot].view(np.int32).reshape(full, T, 4, R, k)
                        wf = wi.view(np.float32)
                        yield (mine[:full], [wi[:, j, 0] for j in range(T)], [wi[:, j, 1] for j in range(T)],
                               [wf[:, j, 2] for j in range(T)], [wf[:, j, 3] for j in range(T)])
                    if ragged:
                        r = self.rows_per_edge * (self.n - mine[-1] * self.bs)
                        ti = raw[full * self.slot: full * self.slot + T * 16 * r * k].view(np.int32).reshape(T, 4, r, k)
                        tf = ti.view(np.float32)
                        yield (mine[full:], [ti[None, j, 0] for j in range(T)], [ti[None, j, 1] for j in range(T)],
                               [tf[None, j, 2] for j in range(T)], [tf[None, j, 3] for j in range(T)])
                    continue
                for q, b in enumerate(mine):
                    edges = min(self.bs, self.n - b * self.bs)
                    r = self.rows_per_edge * edges
                    ti = raw[q * self.slot: q * self.slot + T * 16 * r * k].view(np.int32).reshape(T, 4, r, k)
                    tf = ti.view(np.float32)
                    yield (b, [ti[j, 0] for j in range(T)], [ti[j, 1] for j in range(T)], [tf[j, 2] for j in range(T)], [tf[j, 3] for j in range(T)])
        finally:
            self.close()

    def close(self):
        if getattr(self, "_ep", None):
            self.f._L.zebra_tppr_epoch_end(self._ep)
            self._ep = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
