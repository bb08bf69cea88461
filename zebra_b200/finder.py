"""Drop-in replacement for the reference's numba `tppr_finder` (LuckyLYM/Zebra utils/util.py:391-873).

Same constructor, attributes, method names, argument meaning and return types; the per-node score
dictionaries are device-resident fixed-capacity top-k tables and every update runs in the
sm_100a span kernel behind the C ABI (include/zebra_tppr.h).  There is no CPU path: constructing
a finder without a CUDA device raises.

Snapshot semantics.  In the reference `backup_tppr`, `restore_val_tppr` and the cache written by
`compute_val_tppr` are shallow numba typed-List copies, i.e. they ALIAS the live containers
(utils/util.py:436-444, 872-873).  The default here reproduces exactly that by
handing out references to the same device tables, because parity is judged against what the
reference does.  `ZEBRA_TPPR_SNAPSHOT=strict` (or strict_snapshots=True) switches the three calls
to real device-to-device clones, the semantics the reference's author presumably intended.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref

import numpy as np
import torch

from ._lib import check, lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, torch.Tensor):
        return C.c_void_p(a.data_ptr())
    return a.ctypes.data_as(C.c_void_p)


class tppr_finder:  # noqa: N801  (name kept from the reference)
    _live = []          # weak references to the finders alive in this process, oldest first (batches.install)

    def __init__(self, num_nodes, k, n_tppr, alpha_list, beta_list, device=None, strict_snapshots=None, tie_mode=None):
        if not torch.cuda.is_available():
            raise RuntimeError("zebra_b200.tppr_finder needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.num_nodes = int(num_nodes)
        self.k = int(k)
        self.n_tppr = int(n_tppr)
        self.alpha_list = [float(a) for a in alpha_list]
        self.beta_list = [float(b) for b in beta_list]
        if len(self.alpha_list) != self.n_tppr or len(self.beta_list) != self.n_tppr:
            raise ValueError("alpha_list and beta_list must have n_tppr entries")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if strict_snapshots is None:
            strict_snapshots = os.environ.get("ZEBRA_TPPR_SNAPSHOT", "").lower() == "strict"
        self.strict_snapshots = bool(strict_snapshots)
        self._L = lib()
        h = C.c_void_p()
        a = (C.c_double * self.n_tppr)(*self.alpha_list)
        b = (C.c_double * self.n_tppr)(*self.beta_list)
        check(self._L.zebra_tppr_create(C.byref(h), self.num_nodes, self.k, self.n_tppr, a, b, self.device.index or 0))
        self._h = h
        if tie_mode is None:
            tie_mode = os.environ.get("ZEBRA_TIE_MODE", "exact")
        self.tie_mode = {"exact": 0, "stable": 1, 0: 0, 1: 1}[tie_mode]   # include/zebra_tppr.h zebra_tppr_set_tie_mode
        if self.tie_mode:
            check(self._L.zebra_tppr_set_tie_mode(h, self.tie_mode))
        if os.environ.get("ZEBRA_MAX_BLOCKS"):            # developer knob: cap the persistent grid (blocks of 8 warps)
            check(self._L.zebra_tppr_set_max_blocks(h, int(os.environ["ZEBRA_MAX_BLOCKS"])))
        self._table_bytes = self._L.zebra_tppr_table_bytes(h)
        self._record_bytes = self._L.zebra_tppr_record_bytes(h)
        self._host_out = None
        self._tables = None
        self._sched = None              # an announced pass (schedule_pass), see _ScheduledPass
        tppr_finder._live[:] = [r for r in tppr_finder._live if r() is not None] + [weakref.ref(self)]
        self.reset_val_tppr()
        self.reset_tppr()

    # ---- state management (utils/util.py:402-444) ----
    def _new_tables(self):
        return [torch.zeros(self._table_bytes, dtype=torch.uint8, device=self.device) for _ in range(self.n_tppr)]

    def _bind(self):
        arr = (C.c_void_p * self.n_tppr)(*[t.data_ptr() for t in self._tables])
        check(self._L.zebra_tppr_bind_tables(self._h, arr))

    def _snapshot(self, tables):
        return [t.clone() for t in tables] if self.strict_snapshots else list(tables)

    def reset_tppr(self):
        self._settle()
        self._tables = self._new_tables()
        self._bind()

    def reset_val_tppr(self):
        self._val_tables = self._new_tables()

    def backup_tppr(self):
        self._settle()
        return self._snapshot(self._tables)

    def restore_tppr(self, backup):
        self._settle()
        self._tables = list(backup)
        self._bind()

    def restore_val_tppr(self):
        self._settle()
        self._tables = self._snapshot(self._val_tables)
        self._bind()

    # ---- streaming (utils/util.py:473-576, 581-679, 682-782, 787-873) ----
    def _stream_ptr(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _launch(self, source_nodes, timestamps, edge_idxs, rows_per_edge, inst, to_host):
        nodes = np.ascontiguousarray(source_nodes, dtype=np.int32)
        ts = np.ascontiguousarray(timestamps, dtype=np.float64)
        eidx = np.ascontiguousarray(edge_idxs, dtype=np.int64)
        n = len(eidx)
        rows = rows_per_edge * n
        if len(nodes) != max(rows_per_edge, 2) * n or len(ts) < n:
            raise ValueError("source_nodes/timestamps do not match edge_idxs")
        nbytes = self._L.zebra_tppr_out_bytes(self._h, n, rows_per_edge)
        host = None
        if self._sched is not None:
            # an announced pass: the rows of this batch were computed with its window, ahead of the call
            hit = self._sched.take(nodes, ts, eidx, rows_per_edge, inst)
            if hit is not None:
                if to_host and nbytes:
                    if self._host_out is None or self._host_out.numel() < nbytes:
                        self._host_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
                        self._dev_scratch = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
                    host = self._host_out
                    host[:nbytes].copy_(hit[:nbytes])          # synchronous D2H into the pinned buffer
                return hit, host, rows, nbytes
        if to_host and nbytes:
            # numpy-returning calls: the device copy of the rows is scratch, kept across calls (no allocation per batch)
            if self._host_out is None or self._host_out.numel() < nbytes:
                self._host_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
                self._dev_scratch = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            host, dev_out = self._host_out, self._dev_scratch
        else:
            dev_out = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            check(self._L.zebra_tppr_stream(self._h, _ptr(nodes), _ptr(ts), _ptr(eidx), n, rows_per_edge, inst,
                                            _ptr(dev_out) if nbytes else None, _ptr(host), self._stream_ptr()))
        return dev_out, host, rows, nbytes

    def streaming_topk_raw(self, source_nodes, timestamps, edge_idxs, rows_per_edge=3):
        """The span kernel's output left on the GPU as one uint8 buffer laid out [n_tppr][4][rows*k]
        (int32 node | int32 eidx | f32 dt | f32 w): the input of the fused gather.  Asynchronous."""
        dev_out, _, rows, _ = self._launch(source_nodes, timestamps, edge_idxs, rows_per_edge, -1, False)
        return dev_out, rows

    def _run(self, source_nodes, timestamps, edge_idxs, rows_per_edge, inst, to_host):
        dev_out, host, rows, nbytes = self._launch(source_nodes, timestamps, edge_idxs, rows_per_edge, inst, to_host)
        plane = rows * self.k
        if to_host:
            # one copy out of the pinned buffer (it is overwritten by the next call), then views of it
            buf = host.numpy()[:nbytes].copy() if nbytes else np.zeros(0, np.uint8)
            outs = ([], [], [], [])
            for j in range(self.n_tppr):
                base = j * plane * 16
                for q, dt in enumerate((np.int32, np.int32, np.float32, np.float32)):
                    outs[q].append(buf[base + q * plane * 4: base + (q + 1) * plane * 4].view(dt).reshape(rows, self.k))
            return outs
        outs = ([], [], [], [])
        for j in range(self.n_tppr):
            base = j * plane * 16
            for q, dt in enumerate((torch.int32, torch.int32, torch.float32, torch.float32)):
                outs[q].append(dev_out[base + q * plane * 4: base + (q + 1) * plane * 4].view(dt).view(rows, self.k))
        return outs

    def streaming_topk(self, source_nodes, timestamps, edge_idxs):
        """[src;dst;neg] int32[3n], f64[3n], int64[n] -> 4 lists of n_tppr arrays [3n,k] (i32,i32,f32,f32)."""
        return self._run(source_nodes, timestamps, edge_idxs, 3, -1, True)

    def streaming_topk_no_fake(self, source_nodes, timestamps, edge_idxs):
        return self._run(source_nodes, timestamps, edge_idxs, 2, -1, True)

    def single_streaming_topk(self, source_nodes, timestamps, edge_idxs, tppr_id):
        o = self._run(source_nodes, timestamps, edge_idxs, 3, int(tppr_id), True)
        return tuple(x[int(tppr_id)] for x in o)

    def streaming_topk_device(self, source_nodes, timestamps, edge_idxs, rows_per_edge=3):
        """Level-1 fast path: same result as streaming_topk, left on the GPU as torch tensors
        (asynchronous on the current stream; nothing is copied back)."""
        return self._run(source_nodes, timestamps, edge_idxs, rows_per_edge, -1, False)

    def compute_val_tppr(self, sources, targets, timestamps, edge_idxs):
        src = np.ascontiguousarray(sources, dtype=np.int64)
        dst = np.ascontiguousarray(targets, dtype=np.int64)
        ts = np.ascontiguousarray(timestamps, dtype=np.float64)
        eidx = np.ascontiguousarray(edge_idxs, dtype=np.int64)
        self._settle()
        self._update_only(src, dst, ts, eidx)
        self._val_tables = self._snapshot(self._tables)

    def _update_only(self, src, dst, ts, eidx):
        src = np.ascontiguousarray(src, dtype=np.int64)
        dst = np.ascontiguousarray(dst, dtype=np.int64)
        with torch.cuda.device(self.device):
            check(self._L.zebra_tppr_update_only(self._h, _ptr(src), _ptr(dst), _ptr(ts), _ptr(eidx), len(src),
                                                 self._stream_ptr()))

    # ---- pipelined epochs (include/zebra_tppr.h "pipelined epochs") ----
    def stream_epoch(self, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge=3, window=128,
                     windows=False):
        """Generator over the batches of a chronological span held in HOST memory.

        Yields, batch after batch, exactly what `streaming_topk(cat[src_b,dst_b,neg_b], ts_b, eidx_b)`
        (rows_per_edge=3) or `streaming_topk_no_fake` (2) would return for that batch — four lists of
        n_tppr numpy arrays [rows,k] — while the library keeps later windows of `window` batches in
        flight (host dependency pass, H2D, span kernel, D2H into pinned memory).  The arrays are views
        of pinned library memory, valid until the generator is advanced past the end of their window
        (copy them to keep them).  rows_per_edge=0 updates only and yields nothing (compute_val_tppr
        semantics without the snapshot).  `windows=True` yields one (first_batch, nodes, eidx, dt, w)
        tuple per window instead, each a list of n_tppr arrays [n_batches, rows, k] (full batches only;
        a ragged last batch is yielded as its own window)."""
        src = np.ascontiguousarray(src, dtype=np.int32)
        dst = np.ascontiguousarray(dst, dtype=np.int32)
        neg = None if neg is None else np.ascontiguousarray(neg, dtype=np.int32)
        ts = np.ascontiguousarray(timestamps, dtype=np.float64)
        eidx = np.ascontiguousarray(edge_idxs, dtype=np.int64)
        n, bs, k, T = len(src), int(batch_size), self.k, self.n_tppr
        if not (len(dst) == n and len(eidx) == n and len(ts) >= n and (neg is None or len(neg) == n)):
            raise ValueError("src/dst/neg/timestamps/edge_idxs do not match")
        if n == 0:
            return
        self._settle()
        L = self._L
        ep = C.c_void_p()
        with torch.cuda.device(self.device):
            check(L.zebra_tppr_epoch_begin(self._h, _ptr(src), _ptr(dst), _ptr(neg), _ptr(ts), _ptr(eidx), n, rows_per_edge,
                                           bs, int(window), self._stream_ptr(), C.byref(ep)))
        try:
            slot = int(L.zebra_tppr_epoch_slot_bytes(ep))
            rows_ptr, b0, nb = C.c_void_p(), C.c_int64(), C.c_int64()
            R = rows_per_edge * bs
            while True:
                check(L.zebra_tppr_epoch_next(ep, C.byref(rows_ptr), C.byref(b0), C.byref(nb)))
                if nb.value == 0:
                    break
                if not rows_per_edge:
                    continue
                first, cnt = b0.value, nb.value
                raw = np.ctypeslib.as_array(C.cast(rows_ptr, C.POINTER(C.c_uint8)), shape=(cnt * slot,))
                last_edges = min(bs, n - (first + cnt - 1) * bs)
                full = cnt if last_edges == bs else cnt - 1
                if full:
                    wi = raw[:full * slot].view(np.int32).reshape(full, T, 4, R, k)
                    wf = wi.view(np.float32)
                    if windows:
                        yield (first, [wi[:, j, 0] for j in range(T)], [wi[:, j, 1] for j in range(T)],
                               [wf[:, j, 2] for j in range(T)], [wf[:, j, 3] for j in range(T)])
                    else:
                        for q in range(full):
                            bi, bf = wi[q], wf[q]
                            yield ([bi[j, 0] for j in range(T)], [bi[j, 1] for j in range(T)],
                                   [bf[j, 2] for j in range(T)], [bf[j, 3] for j in range(T)])
                if full < cnt:       # ragged last batch of the span: same layout with fewer rows
                    r = rows_per_edge * last_edges
                    ti = raw[full * slot: full * slot + T * 16 * r * k].view(np.int32).reshape(T, 4, r, k)
                    tf = ti.view(np.float32)
                    if windows:
                        yield (first + full, [ti[None, j, 0] for j in range(T)], [ti[None, j, 1] for j in range(T)],
                               [tf[None, j, 2] for j in range(T)], [tf[None, j, 3] for j in range(T)])
                    else:
                        yield ([ti[j, 0] for j in range(T)], [ti[j, 1] for j in range(T)],
                               [tf[j, 2] for j in range(T)], [tf[j, 3] for j in range(T)])
        finally:
            L.zebra_tppr_epoch_end(ep)

    # ---- announced passes (the batches of a pass are known before its loop) ----
    def schedule_pass(self, src, dst, neg, timestamps, edge_idxs, batch_size, window=64):
        """Announce the batches the next `streaming_topk*` calls will bring (a validation or test pass:
        train.py:196-203 slices them one by one, evaluation/evaluation.py:19-33 draws their seeded negatives one by
        one — both are known before the loop starts).  The span goes to the device once (a StreamPlan) and is
        streamed `window` batches per launch AHEAD of the calls; a call whose arguments equal the next announced
        batch returns the rows already computed for it (same bytes, same final state as per-call streaming).
        A call that brings anything else, or any state call (reset / backup / restore / ...), ends the
        announcement first: the tables are put back to the start of the window in flight, the batches consumed so
        far are replayed, and the call proceeds as usual — announcing a pass never changes a result."""
        self._settle()
        sp = _ScheduledPass(self, src, dst, neg, timestamps, edge_idxs, batch_size, window)
        if sp.n_batches:
            self._sched = sp
        return sp

    def _settle(self):
        if self._sched is not None:
            self._sched.abandon()

    # ---- resident spans (include/zebra_tppr.h "plans") ----
    def make_plan(self, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge=3):
        """Prepare a whole chronological span once and keep it resident in HBM."""
        return StreamPlan(self, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge)

    # ---- introspection (tests) ----
    def state(self, inst, nodes):
        """Vectors of `nodes` in instance `inst`: list of (eidx, node, ts, score, norm) in slot order."""
        self._settle()          # an announced pass runs ahead of its calls: report the state the calls have reached
        ids = np.ascontiguousarray(nodes, dtype=np.int64)
        m, k = len(ids), self.k
        cnt = np.zeros(m, np.int32); e = np.zeros((m, k), np.int32); nd = np.zeros((m, k), np.int32)
        t = np.zeros((m, k), np.float64); s = np.zeros((m, k), np.float64); nm = np.zeros(m, np.float64)
        with torch.cuda.device(self.device):
            check(self._L.zebra_tppr_read_state(self._h, int(inst), _ptr(ids), m, _ptr(cnt), _ptr(e), _ptr(nd), _ptr(t),
                                                _ptr(s), _ptr(nm), self._stream_ptr()))
        return [(e[a, :cnt[a]].astype(np.int64), nd[a, :cnt[a]].astype(np.int64), t[a, :cnt[a]], s[a, :cnt[a]], float(nm[a]))
                for a in range(m)]

    def sync_check(self):
        check(self._L.zebra_tppr_sync_check(self._h, self._stream_ptr()))

    def counters(self):
        o = np.zeros(8, np.int64)
        check(self._L.zebra_tppr_counters(self._h, _ptr(o)))
        return {"items": int(o[0]), "tie_prunes": int(o[1]), "launches": int(o[2]), "timeouts": int(o[3]),
                "merge_ranked": int(o[4]), "chain_items": int(o[5]), "chain_steps": int(o[6])}

    def profile(self):
        o = np.zeros(16, np.uint64)
        check(self._L.zebra_tppr_profile(self._h, _ptr(o)))
        return o

    def last_kernel_ms(self):
        return float(self._L.zebra_tppr_last_kernel_ms(self._h))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._L.zebra_tppr_destroy(self._h)
                self._h = None
        except Exception:
            pass


class StreamPlan:
    """A span resident on the device; `run` replays batches [first, first+n) without host traffic.

    Batch b's rows land in slot (b - first) % out_slots of the returned buffer, each slot laid out
    exactly like one `streaming_topk` call on that batch; `views(buf, b, first)` slices it."""

    def __init__(self, finder, src, dst, neg, timestamps, edge_idxs, batch_size, rows_per_edge):
        self.f = finder
        self.rows_per_edge = int(rows_per_edge)
        self.batch_size = int(batch_size)
        src = np.ascontiguousarray(src, dtype=np.int32)
        dst = np.ascontiguousarray(dst, dtype=np.int32)
        neg = None if neg is None else np.ascontiguousarray(neg, dtype=np.int32)
        ts = np.ascontiguousarray(timestamps, dtype=np.float64)
        eidx = np.ascontiguousarray(edge_idxs, dtype=np.int64)
        self.n = len(src)
        h = C.c_void_p()
        with torch.cuda.device(finder.device):
            check(finder._L.zebra_tppr_plan_create(finder._h, _ptr(src), _ptr(dst), _ptr(neg), _ptr(ts), _ptr(eidx), self.n,
                                                   self.rows_per_edge, self.batch_size, C.byref(h)))
        self._p = h
        self.n_batches = int(finder._L.zebra_tppr_plan_batches(h))
        self.device_bytes = int(finder._L.zebra_tppr_plan_bytes(h))        # what the plan uploaded (H2D) and keeps in HBM
        self.slot_bytes = int(finder._L.zebra_tppr_out_bytes(finder._h, self.batch_size, self.rows_per_edge))

    def alloc_out(self, out_slots):
        return torch.empty(max(1, out_slots * self.slot_bytes), dtype=torch.uint8, device=self.f.device)

    def run(self, first=0, n_batches=None, batches_per_launch=1, out=None, out_slots=None):
        n_batches = self.n_batches - first if n_batches is None else n_batches
        if out_slots is None:
            out_slots = min(batches_per_launch, n_batches)
        if out is None and self.rows_per_edge:
            out = self.alloc_out(out_slots)
        f = self.f
        if f._sched is not None and f._sched.plan is not self:
            f._settle()         # another plan while a pass is announced: that pass ends first
        with torch.cuda.device(f.device):
            check(f._L.zebra_tppr_plan_run(f._h, self._p, first, n_batches, batches_per_launch,
                                           _ptr(out) if self.rows_per_edge else None, out_slots, f._stream_ptr()))
        return out

    def batch_edges(self, b):
        return min(self.batch_size, self.n - b * self.batch_size)

    def views(self, out, b, first=0, out_slots=None):
        """The four lists of n_tppr [rows,k] tensors of batch b inside `out`."""
        out_slots = out.numel() // self.slot_bytes if out_slots is None else out_slots
        slot = (b - first) % out_slots
        nb, k = self.batch_edges(b), self.f.k
        rows = self.rows_per_edge * nb
        plane = rows * k
        base0 = slot * self.slot_bytes
        res = ([], [], [], [])
        for j in range(self.f.n_tppr):
            base = base0 + j * plane * 16
            for q, dt in enumerate((torch.int32, torch.int32, torch.float32, torch.float32)):
                res[q].append(out[base + q * plane * 4: base + (q + 1) * plane * 4].view(dt).view(rows, k))
        return res

    def close(self):
        if getattr(self, "_p", None):
            self.f._L.zebra_tppr_plan_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _ScheduledPass:
    """State of `tppr_finder.schedule_pass`: the plan of the pass, the window computed ahead, the copy of the
    tables taken at the start of that window (what a mismatching call rolls back to)."""

    def __init__(self, finder, src, dst, neg, timestamps, edge_idxs, batch_size, window):
        self.f = finder
        self.bs = int(batch_size)
        self.window = max(1, int(window))
        self.src = np.ascontiguousarray(src, dtype=np.int32)
        self.dst = np.ascontiguousarray(dst, dtype=np.int32)
        self.neg = np.ascontiguousarray(neg, dtype=np.int32)
        self.ts = np.ascontiguousarray(timestamps, dtype=np.float64)
        self.eidx = np.ascontiguousarray(edge_idxs, dtype=np.int64)
        self.n = len(self.src)
        self.n_batches = (self.n + self.bs - 1) // self.bs if self.n else 0
        self.next = 0                   # next batch a call is expected to bring
        self.w0 = self.w1 = 0           # batches [w0, w1) have been streamed
        self.snap = self.snap_ids = None    # the records of the window's endpoints before batch w0
        self.hits = self.launches = 0
        self.rolled_back = False
        self.plan = self.out = self.ends = None
        if self.n_batches:
            self.plan = finder.make_plan(self.src, self.dst, self.neg, self.ts, self.eidx, self.bs, 3)
            self.out = self.plan.alloc_out(min(self.window, self.n_batches))
            # endpoints of every edge on the device, once per pass (index lists of the per-window record copies)
            self.ends = torch.from_numpy(np.stack([self.src, self.dst])).to(finder.device)         # int32 [2, n]

    def _rows(self, table):
        return table.view(torch.int64).view(-1, self.f._record_bytes // 8)

    def _matches(self, b, nodes, ts, eidx):
        lo = b * self.bs
        n = min(self.bs, self.n - lo)
        hi = lo + n
        if len(eidx) != n or len(nodes) != 3 * n or len(ts) < n:
            return False
        # plain elementwise compares: np.array_equal's argument handling costs more than the compare at this size
        return bool((eidx == self.eidx[lo:hi]).all() and (nodes[:n] == self.src[lo:hi]).all()
                    and (nodes[n:2 * n] == self.dst[lo:hi]).all() and (nodes[2 * n:] == self.neg[lo:hi]).all()
                    and (ts[:n] == self.ts[lo:hi]).all())

    def take(self, nodes, ts, eidx, rows_per_edge, inst):
        """The device rows of the call if it brings the next announced batch, else None (after settling)."""
        b = self.next
        if rows_per_edge != 3 or inst != -1 or b >= self.n_batches or not self._matches(b, nodes, ts, eidx):
            self.abandon()
            return None
        f = self.f
        if b >= self.w1:
            cnt = min(self.window, self.n_batches - b)
            # what a roll-back needs: the records of the window's endpoints as they are now (negatives are only
            # read).  Duplicates stay in the index list: every copy of a node carries the same record.
            lo, hi = b * self.bs, min(self.n, (b + cnt) * self.bs)
            self.snap_ids = self.ends[:, lo:hi].reshape(-1)
            self.snap = [self._rows(t).index_select(0, self.snap_ids) for t in f._tables]
            self.plan.run(b, cnt, cnt, out=self.out, out_slots=cnt)
            self.w0, self.w1 = b, b + cnt
            self.launches += 1
        slot = b - self.w0
        nb = self.plan.batch_edges(b)
        nbytes = f._L.zebra_tppr_out_bytes(f._h, nb, 3)
        rows = self.out[slot * self.plan.slot_bytes: slot * self.plan.slot_bytes + max(nbytes, 1)]
        self.next = b + 1
        self.hits += 1
        if self.next == self.n_batches:
            self._finish()
        return rows

    def abandon(self):
        """End the announcement: leave the tables as per-call streaming of the consumed batches would."""
        f = self.f
        if f._sched is self:
            f._sched = None
        if self.plan is None:
            return
        redo = None
        if self.next < self.w1:
            # the window ran ahead of the calls: back to its start, then the consumed part of it again (as one
            # update-only span: a plan's batches run strictly in order, and rows are not needed a second time)
            for t, s0 in zip(f._tables, self.snap):
                self._rows(t).index_copy_(0, self.snap_ids.long(), s0)
            if self.next > self.w0:
                lo, hi = self.w0 * self.bs, self.next * self.bs
                redo = (self.src[lo:hi], self.dst[lo:hi], self.ts[lo:hi], self.eidx[lo:hi])
            self.rolled_back = True
        self._finish()
        if redo is not None:
            f._update_only(*redo)

    def _finish(self):
        f = self.f
        if f._sched is self:
            f._sched = None
        self.snap = self.snap_ids = None
        if self.plan is not None:
            torch.cuda.current_stream(f.device).synchronize()    # the plan's device arrays are freed below
            self.plan.close()
            self.plan = None
        # what the pass held: the ring stays alive only through the rows handed out last
        self.out = self.ends = self.src = self.dst = self.neg = self.ts = self.eidx = None

    cancel = abandon
