"""bench.py — T-PPR maintenance throughput on B200 (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

Workload (config.workload): by default BASELINE.json configs[4], the stream the metric's target is
quoted on — Wiki-Talk-shaped synthetic unipartite stream (1,140,149 nodes, 7,833,140 edges), streaming
T-PPR with negatives, topk=20, two (alpha,beta) pairs, batches of 200 edges.  `--workload` selects the
shape of another BASELINE config (wikipedia, reddit, mooc, superuser) with the same line format.  One
step = reset the tables and stream the whole span once.

  value   edges/s with the span resident in HBM (a "plan": edge records + dependency tickets uploaded
          once), outputs written to HBM, timed with CUDA events on the launching stream.
  e2e     edges/s through tppr_finder.stream_epoch (host numpy in, host numpy out): the per-batch
          streaming_topk results of the whole span; host dependency pass, host->device copies, kernels and
          the device->host copy of every batch's 4*n_tppr result arrays are inside the timed region.
          The strict one-call-per-batch drop-in is reported beside it (e2e.per_call_drop_in).
  roofline / cpu_baseline: see DESIGN.md.  The CPU leg executes oracle/ (the C port of the reference's
          single-threaded numba path) as a reported baseline only.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K, BS = 20, 200
ALPHAS, BETAS = [0.1, 0.1], [0.5, 0.95]
WORKLOAD = "wikitalk"
WORKLOADS = ("wikipedia", "reddit", "mooc", "superuser", "wikitalk")      # BASELINE.json configs[0..4] shapes
METRIC, UNIT = "tppr_streaming_edges_per_sec", "edges/s"


def bytes_per_edge(k=K, n_tppr=len(ALPHAS)):
    """Algorithmic bytes of one streamed edge: per instance read 3 vectors, write 2,
    48k output bytes, 28-byte edge record.  V(k) = 24k + 8."""
    return n_tppr * (5 * (24 * k + 8) + 48 * k + 28)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per span-kernel launch (read + write) from the newest committed `ncu --set full` capture
    of this command (profiles/*ncu_full_span*_raw.csv), or None."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*ncu_full_span*_raw.csv")))
    if not files:
        return None, None
    try:
        rows = list(csv.reader(open(files[-1])))
        hdr, units, data = rows[0], rows[1], rows[2:]
        tot = []
        for r in data:
            b = 0.0
            for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                i = hdr.index(name)
                scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
                b += float(r[i]) * scale
            tot.append(b)
        return (sum(tot) / len(tot), os.path.relpath(files[-1], ROOT)) if tot else (None, None)
    except Exception:
        return None, None


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_workload(n_edges=None, seed=None):
    from zebra_b200 import synth
    st = synth.make_stream(WORKLOAD, n_edges=n_edges, seed=seed)
    neg = synth.make_negatives(st, BS)
    return st, neg


def cpu_port_throughput(st, neg, sample_edges, threads):
    """The oracle (C port of the reference's numba path) on the first `sample_edges` edges of the
    workload, bs=200 batches with negatives.  threads > 1 runs the independent (alpha,beta) instances
    concurrently (ctypes releases the GIL); the reference itself is single-threaded."""
    import oracle
    n = min(sample_edges, st.n_edges)
    N = st.shape.table_nodes
    finders = ([oracle.OracleFinder(N, K, 1, [a], [b]) for a, b in zip(ALPHAS, BETAS)] if threads > 1
               else [oracle.OracleFinder(N, K, len(ALPHAS), ALPHAS, BETAS)])
    batches = []
    for lo in range(0, n, BS):
        hi = min(n, lo + BS)
        batches.append((np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32),
                        np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi]))

    def work(f):
        for b in batches:
            f.streaming_topk(*b)

    t = time.perf_counter()
    if len(finders) == 1:
        work(finders[0])
    else:
        ths = [threading.Thread(target=work, args=(f,)) for f in finders]
        [x.start() for x in ths]
        [x.join() for x in ths]
    dt = time.perf_counter() - t
    return n / dt, n, dt


_NUMBA_SNIPPET = r"""
import json, os, sys, time
import numpy as np
ref, root, name, n, bs = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
sys.dont_write_bytecode = True
sys.path.insert(0, root)
from zebra_b200 import synth
st = synth.make_stream(name, n_edges=n)
neg = synth.make_negatives(st, bs)
sys.path.insert(0, ref)
t0 = time.perf_counter()
from utils.util import tppr_finder                      # the reference's own numba jitclass, unmodified
f = tppr_finder(st.shape.table_nodes, 20, 2, [0.1, 0.1], [0.5, 0.95])
def batch(lo, hi):
    return (np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32),
            np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi].astype(np.int64))
f.streaming_topk(*batch(0, bs))                        # compiles this argument-type signature
compile_s = time.perf_counter() - t0
f.reset_tppr()
t = time.perf_counter()
for lo in range(0, n, bs):
    f.streaming_topk(*batch(lo, min(n, lo + bs)))
dt = time.perf_counter() - t
print(json.dumps({"value": n / dt, "edges": n, "seconds": dt, "compile_s": compile_s}))
"""


def numba_reference_throughput(sample_edges, timeout_s=240):
    """The reference ITSELF (LuckyLYM/Zebra utils/util.py:391-576, numba jitclass, single-threaded by
    construction) on a prefix of the workload: imported unmodified from baseline/_ref (a copy of the
    reference tree made by __graft_entry__.build()), compiled once, then timed with perf_counter around
    the streaming_topk calls only.  Runs in a subprocess so that a missing numba
    or a slow compile cannot take the bench down; returns a dict for the JSON line."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref, "utils", "util.py")):
        return {"unavailable": "baseline/_ref missing (made by __graft_entry__.build() where /root/reference exists)"}
    try:
        r = subprocess.run([sys.executable, "-c", _NUMBA_SNIPPET, ref, ROOT, WORKLOAD, str(sample_edges), str(BS)],
                           capture_output=True, text=True, timeout=timeout_s,
                           env=dict(os.environ, PYTHONDONTWRITEBYTECODE="1", NUMBA_NUM_THREADS="1"))
        if r.returncode != 0:
            return {"unavailable": (r.stderr.strip().splitlines() or ["numba run failed"])[-1][:200]}
        d = json.loads(r.stdout.strip().splitlines()[-1])
        return {"value": d["value"], "unit": UNIT, "cores": f"1 of {os.cpu_count()}", "kind": "reference",
                "sample": f"first {d['edges']} edges of the workload, bs={BS}, with negatives, two (alpha,beta) instances "
                          f"({d['seconds']:.1f} s timed after a {d['compile_s']:.0f} s import + numba compile)"}
    except subprocess.TimeoutExpired:
        return {"unavailable": f"numba reference did not finish within {timeout_s} s"}
    except Exception as ex:  # noqa: BLE001
        return {"unavailable": f"{type(ex).__name__}: {ex}"[:200]}


def workload_text(st, world=1):
    if world > 1:
        return (f"{world} {WORKLOAD}-shaped communities merged into one stream, streaming T-PPR topk={K}, alpha {ALPHAS}, "
                f"beta {BETAS}, bs={BS}, negatives included")
    return (f"{WORKLOAD}-shaped synthetic stream ({st.shape.n_nodes} nodes, {st.shape.n_edges} edges), streaming T-PPR "
            f"topk={K}, alpha {ALPHAS}, beta {BETAS}, bs={BS}, negatives included")


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the box's host cores.
    The line's `value` is oracle/ (the C port of the reference's single-threaded numba path, validated
    bit-for-bit against recordings of the reference; one thread per (alpha,beta) instance — the only
    parallelism the algorithm admits), because the numba original needs ~10 minutes for one pass over
    this workload; the numba original itself is timed on a 50 K-edge prefix of the same workload and
    reported next to it (cpu_baseline.reference_numba)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = len(ALPHAS)          # the only parallelism the algorithm admits: independent instances
    sample = args.ref_sample_edges
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        # same workload as our arm at this N: `world` communities merged into one stream
        from zebra_b200 import synth
        st, neg = synth.make_communities(WORKLOAD, world, BS, n_edges=-(-sample // world))
        full = synth.SHAPES[WORKLOAD]
    else:
        from zebra_b200 import synth
        full = synth.SHAPES[WORKLOAD]
        st, neg = load_workload(n_edges=min(sample, full.n_edges))
    sample = min(sample, st.n_edges)
    for _ in range(max(1, min(args.warmup, 1))):
        cpu_port_throughput(st, neg, min(sample, 100_000), threads)
    vals, t_all = [], 0.0
    for _ in range(args.steps):
        v, n, dt = cpu_port_throughput(st, neg, sample, threads)
        vals.append(v); t_all += dt
    value = float(np.mean(vals))
    numba = numba_reference_throughput(args.numba_sample_edges) if not args.no_numba else {"unavailable": "--no-numba"}
    import dataclasses
    st_full = dataclasses.replace(st, shape=full) if world == 1 else st
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_all / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_text(st_full, world),
                   "sample": f"each step streams the first {sample} edges of that workload (the C port needs ~26 s, the numba "
                             f"original ~10 min for all of it)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "of_cores": os.cpu_count(), "kind": "port",
                         "sample": f"first {sample} edges of the workload, bs={BS}, with negatives; one thread per (alpha,beta) instance",
                         "reference_numba": numba},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_sharded(args, torch, dist, world, rank, local, dev):
    """N > 1: T-PPR state partitioned by node id over the N GPUs (zebra_b200/sharded.py).  Every rank
    walks the same replicated edge stream; the span kernel pushes cross-shard old vectors through
    peer-mapped mailboxes over NVLink and stores every output row into every rank's output region.

    Weak scaling (the `value`): the graph grows with N — N wikitalk-shaped communities (N x nodes,
    N x edges, hub degrees of one community), each spread over all ranks, so (N-1)/N of the edges
    cross shards.  Strong scaling (extra key `strong`): the N=1 stream itself sharded over N GPUs;
    its hottest node's dependency chain is as long as on one GPU, so it cannot speed up."""
    from zebra_b200 import synth
    from zebra_b200.sharded import create_sharded, exchange_stats
    # batches per launch grow with N so that one launch holds as much work per community as at N=1
    # (every launch boundary drains all dependency chains on all ranks)
    window = args.shard_batches_per_launch * world
    slot = len(ALPHAS) * 16 * 3 * BS * K
    main = torch.cuda.current_stream()
    copy_stream = torch.cuda.Stream()

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def parity_check(f, plan, st, neg):
        """Before anything is timed: the rows THIS rank is delivered for the first batches of the stream
        (its batches b % N == rank, stored by their owners over NVLink) against the oracle, bit for bit.
        The oracle runs on node ids compacted to the prefix (its tables would need GBs otherwise)."""
        import oracle
        nb = min(plan.n_batches, args.parity_batches)
        if nb <= 0:
            return {"edges": 0, "ok": True}
        n = min(st.n_edges, nb * BS)
        uniq = np.unique(np.concatenate([st.src[:n], st.dst[:n], neg[:n]]))
        back = np.concatenate([[0], uniq]).astype(np.int32)                 # compact id -> node id (0 = padding)
        comp = lambda a: (np.searchsorted(uniq, a) + 1).astype(np.int32)   # noqa: E731
        cs, cd, cn = comp(st.src[:n]), comp(st.dst[:n]), comp(neg[:n])
        o = oracle.OracleFinder(len(uniq) + 1, K, len(ALPHAS), ALPHAS, BETAS)
        f.reset_tppr()
        barrier()
        ok, checked = True, 0
        for b0 in range(0, nb, window):
            cnt_b = min(window, nb - b0)
            plan.run(b0, cnt_b, window, out_slots=window)
            barrier()                                     # every owner's stores have landed
            host = f.out[: cnt_b * slot].cpu().numpy()
            for b in range(b0, b0 + cnt_b):
                lo, hi = b * BS, min(n, (b + 1) * BS)
                ref = o.streaming_topk(np.concatenate([cs[lo:hi], cd[lo:hi], cn[lo:hi]]), np.concatenate([st.ts[lo:hi]] * 3),
                                       st.eidx[lo:hi])
                if b % world != rank:
                    continue
                rows = 3 * (hi - lo)
                buf = host[(b - b0) * slot: (b - b0) * slot + len(ALPHAS) * 16 * rows * K].view(np.int32).reshape(len(ALPHAS), 4, rows, K)
                for j in range(len(ALPHAS)):
                    ok &= bool(np.array_equal(buf[j, 0], back[ref[0][j]]) and np.array_equal(buf[j, 1], ref[1][j])
                               and np.array_equal(buf[j, 2].view(np.float32), ref[2][j]) and np.array_equal(buf[j, 3].view(np.float32), ref[3][j]))
                checked += hi - lo
            barrier()                                     # nobody overwrites the slots before everybody has read
        f.sync_check()
        t = torch.tensor([1.0 if ok else 0.0, float(checked)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        all_ok = bool(t[0] > 0.5)
        if not all_ok:
            raise SystemExit(f"rank {rank}: sharded rows differ from the oracle in the first {n} edges (parity check before timing)")
        return {"edges": n, "batches": nb, "ok": True, "rows_checked_per_rank_min": int(t[1]) * 3,
                "what": "every rank compared the rows delivered to it (batches b % N == rank) with the oracle, bit for bit, before timing"}

    def measure(st, neg, sampler=None, e2e=True):
        E, N = st.n_edges, st.shape.table_nodes
        # rows of batch b are delivered to rank b % N only: each data-parallel rank consumes its own batches
        f = create_sharded(N, K, len(ALPHAS), ALPHAS, BETAS, window * BS, 2 * window * slot, row_mode=1)
        plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, BS)
        n_win = -(-plan.n_batches // window)
        parity = parity_check(f, plan, st, neg)

        def step():
            f.reset_tppr()                       # no barrier needed: records are touched by their owner only
            plan.run(0, plan.n_batches, window, out_slots=window)

        for _ in range(args.warmup):
            step()
        f.sync_check()
        if sampler is not None:
            sampler.start()
        barrier()
        l0 = f.counters()["launches"]
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(args.steps):
            step()
        ev[1].record()
        barrier()
        f.sync_check()
        ms = ev[0].elapsed_time(ev[1])
        launches = f.counters()["launches"] - l0
        res = {"E": E, "N": N, "ms": ms, "launches": launches, "n_launch": n_win, "n_batches": plan.n_batches,
               "record_bytes": f._record_bytes, "parity": parity}
        pb = torch.tensor([float(plan.device_bytes)], dtype=torch.float64, device=dev)
        dist.all_reduce(pb)                       # H2D of a plan, summed over the ranks
        res["plan_bytes"] = int(pb[0])
        if e2e:
            # end to end: host arrays in (replicated span on every rank), this rank's batches out in pinned host
            # memory (rank r consumes batches b % N == r: together the ranks read every batch exactly once).
            # ShardedFinder.epoch = the library's pipelined sharded epoch: per window the ownership-aware
            # dependency pass (worker threads), H2D of the records the rank takes part in, the sharded span
            # kernel, and — once every rank has finished the window (in-library hand-shake through the arena
            # header, no collective) — the D2H of the rank's batches on a second stream.
            def e2e_pass():
                f.reset_tppr()
                acc, cnt_b = 0.0, 0
                for bs_idx, nd, ei, dt, w in f.epoch(st.src, st.dst, neg, st.ts, st.eidx, BS, window=window, windows=True):
                    acc += float(w[-1][:, 0, 0].sum()) + float(nd[0][:, -1, -1].sum())      # every batch's arrays are touched
                    cnt_b += len(bs_idx)
                assert cnt_b == len(range(rank, plan.n_batches, world)), (cnt_b, plan.n_batches)
                return acc

            e2e_pass()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                e2e_pass()
            barrier()
            res["e2e_s"] = time.perf_counter() - t0
            res["e2e_h2d"] = 0
        t = torch.tensor([res["ms"], res.get("e2e_s", 0.0)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res["ms"], res["e2e_s"] = float(t[0]), float(t[1])
        res["tie_prunes"] = f.counters()["tie_prunes"]
        res["chain_steps"] = f.counters()["chain_steps"]
        plan.close()
        f.disconnect()
        barrier()
        f.close()
        return res

    sampler = ClockSampler(local)
    st, neg = synth.make_communities(WORKLOAD, world, BS, n_edges=args.edges)
    xs = exchange_stats(st.src, st.dst, neg, world, K, len(ALPHAS), row_mode=1)
    weak = measure(st, neg, sampler if rank == 0 else None)
    clocks = sampler.stop() if rank == 0 else None
    del st, neg
    strong = None
    if not args.no_strong:
        st1, neg1 = load_workload(n_edges=args.edges)
        strong = measure(st1, neg1, e2e=False)
        strong["exchange"] = exchange_stats(st1.src, st1.dst, neg1, world, K, len(ALPHAS), row_mode=1)
        # the same stream on ONE GPU (rank 0, single-GPU kernel with chain warps), measured in this very run
        if rank == 0:
            from zebra_b200.finder import tppr_finder
            f1 = tppr_finder(st1.shape.table_nodes, K, len(ALPHAS), ALPHAS, BETAS, device=dev)
            p1 = f1.make_plan(st1.src, st1.dst, neg1, st1.ts, st1.eidx, BS)
            w1 = args.batches_per_launch
            o1 = p1.alloc_out(w1)
            best = 1e30
            for _ in range(2):
                f1.reset_tppr()
                p1.run(0, p1.n_batches, w1, out=o1, out_slots=w1)
                f1.sync_check()
                best = min(best, f1.last_kernel_ms())
            strong["single_gpu_ms"] = best
            p1.close()
            del f1, o1
        barrier()
    if rank != 0:
        return
    peak, peak_src = peaks()
    E, ms = weak["E"], weak["ms"]
    value = E * args.steps / (ms * 1e-3)
    per_launch_ms = ms / max(weak["launches"], 1)
    achieved = E * bytes_per_edge() / weak["n_launch"] / (per_launch_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{world} {WORKLOAD}-shaped communities merged into one stream ({weak['N'] - 1} nodes, {E} edges; per GPU "
                               f"{(weak['N'] - 1) // world} nodes of state), streaming T-PPR topk={K}, alpha {ALPHAS}, beta {BETAS}, bs={BS}, "
                               f"negatives included",
                   "step": "reset tables + stream the whole span once",
                   "batches_per_launch": window,
                   "parallelism": f"node-sharded x{world}: owner = id & {world - 1}; cross-shard old vectors pushed through peer-mapped "
                                  f"mailboxes inside the span kernel (NVLink stores), rows of batch b stored by their owners into rank b % N's output region; "
                                  f"no collective on the data path",
                   "exchange": xs,
                   "l2": f"per-GPU tables {len(ALPHAS) * ((weak['N'] + world - 1) // world) * weak['record_bytes'] / 1e6:.0f} MB + "
                         f"outputs exceed the 126 MB L2 (inputs larger than L2)",
                   "tie_path_prunes_last_step_rank0": weak["tie_prunes"],
                   "chain_and_helper_steps_last_step_rank0": weak["chain_steps"]},
        "e2e": {"value": E * args.steps / weak["e2e_s"], "unit": UNIT,
                "h2d_bytes_per_step": weak["plan_bytes"], "d2h_bytes_per_step": len(ALPHAS) * 16 * 3 * E * K,
                "step": f"ShardedFinder.epoch on every rank (host arrays in): per window of {window} batches the ownership-aware dependency "
                        f"pass on worker threads, H2D of the records of the edges the rank takes part in, the sharded span kernel, an in-library "
                        f"cross-rank completion hand-shake, D2H of the rank's batches (b % {world} == r) into pinned host memory on a second "
                        f"stream (every batch read back once)"},
        "gpu_launches": weak["launches"] * world,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak * world, "unit": "GB/s", "frac": achieved / (peak * world),
                     "traffic": None, "peak_source": peak_src + f" x {world} GPUs", "kernel": "tppr_span_shard_kernel",
                     "bytes_per_edge": bytes_per_edge(), "launch_ms": per_launch_ms},
    }
    if strong is not None:
        sv = strong["E"] * args.steps / (strong["ms"] * 1e-3)
        s1 = strong["E"] / (strong["single_gpu_ms"] * 1e-3)
        line["strong"] = {"what": f"the N=1 workload ({strong['E']} edges, one community) node-sharded over {world} GPUs",
                          "value": sv, "unit": UNIT, "ms_per_step": strong["ms"] / args.steps,
                          "single_gpu_value": s1, "speedup_vs_single_gpu": sv / s1,
                          "target": "north_star: >= 5x at 8 GPUs; bounded by the hottest node's serial chain (DESIGN.md 7)",
                          "parity_check": strong["parity"], "exchange": strong["exchange"]}
    line["parity_check"] = weak["parity"]
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from zebra_b200.finder import tppr_finder

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        if args.mode == "sharded":
            run_sharded(args, torch, dist, world, rank, local, dev)
            dist.destroy_process_group()
            return

    split = world > 1 and args.mode == "instances"
    if split and world != len(ALPHAS):
        raise SystemExit("--mode instances places one (alpha,beta) instance per GPU: run it with --gpus %d" % len(ALPHAS))
    # replicas: every rank streams its own replica of the workload (different seed): weak scaling, no exchange.
    # instances: every rank streams the SAME workload for ONE of the (alpha,beta) instances (utils/util.py:489-492:
    # the instances never interact) — strong scaling over the instances, no exchange either.
    st, neg = load_workload(n_edges=args.edges, seed=None if (rank == 0 or split) else 1005 + 100 * rank)
    E, N = st.n_edges, st.shape.table_nodes
    al, be = ([ALPHAS[rank]], [BETAS[rank]]) if split else (ALPHAS, BETAS)
    f = tppr_finder(N, K, len(al), al, be, device=dev)
    plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, BS)
    window = args.batches_per_launch
    out = plan.alloc_out(window)
    n_launch = (plan.n_batches + window - 1) // window

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        f.reset_tppr()
        plan.run(0, plan.n_batches, window, out=out, out_slots=window)

    for _ in range(args.warmup):
        step()
    f.sync_check()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    l0 = f.counters()["launches"]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(args.steps):
        step()
    ev[1].record()
    barrier()
    f.sync_check()
    ms = ev[0].elapsed_time(ev[1])
    launches = f.counters()["launches"] - l0
    tie_prunes = f.counters()["tie_prunes"]
    clocks = sampler.stop() if rank == 0 else None
    cnt = f.counters()
    deg = np.bincount(np.concatenate([st.src, st.dst]), minlength=N)
    hub_edges = int(deg.max())
    chain_info = {"edges_with_a_chain_endpoint": cnt["chain_items"], "chain_steps": cnt["chain_steps"],
                  "what": "nodes whose edges are walked by a resident warp with the vector in registers (DESIGN.md 4.1)"}
    # the same pass with ties in insertion order (zebra_tppr_set_tie_mode 1, "ties within tolerance"): no replay of
    # numba's argsort; reported beside the exact default
    stable_ms = None
    if not args.no_stable:
        f2 = tppr_finder(N, K, len(al), al, be, device=dev, tie_mode="stable")
        p2 = f2.make_plan(st.src, st.dst, neg, st.ts, st.eidx, BS)
        best = 1e30
        for _ in range(2):
            f2.reset_tppr()
            p2.run(0, p2.n_batches, window, out=out, out_slots=window)
            f2.sync_check()
            best = min(best, f2.last_kernel_ms())
        stable_ms = best
        p2.close()
        del f2

    # ---- e2e: host numpy arrays in, host numpy arrays out, for EVERY batch of the workload ----
    # tppr_finder.stream_epoch = the per-batch streaming_topk results of the whole span, pipelined:
    # the dependency pass on the host, H2D of the edge records, the span kernel and D2H of every
    # batch's 4*n_tppr result arrays all happen inside the timed region.
    def epoch_pass():
        # one [n_batches, rows, k] numpy array per result array, instance and window, in pinned host memory:
        # the per-batch streaming_topk results of the whole span; every batch's arrays are touched
        rows, acc = 0, 0.0
        for first, nodes, eidx_, dt, w in f.stream_epoch(st.src, st.dst, neg, st.ts, st.eidx, BS, window=args.e2e_window, windows=True):
            rows += nodes[0].shape[0] * nodes[0].shape[1]
            acc += float(w[-1][:, 0, 0].sum()) + float(nodes[0][:, -1, -1].sum())
        return rows, acc
    f.reset_tppr()
    epoch_pass()                                    # warm-up: allocates the pinned ring once
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        f.reset_tppr()
        rows, _ = epoch_pass()
        assert rows == 3 * E
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    e2e_edges = E
    h2d = ((E + 7) // 8 * 8) * 48
    d2h = len(al) * 16 * 3 * E * K
    launches_e2e = -(-plan.n_batches // args.e2e_window)

    # informational: the strict per-call drop-in (one synchronous streaming_topk per batch)
    nb_call = min(args.e2e_batches, plan.n_batches)
    batches = []
    for b in range(nb_call):
        lo, hi = b * BS, min(E, (b + 1) * BS)
        batches.append((np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32),
                        np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi]))
    f.reset_tppr()
    for b in batches[:50]:
        f.streaming_topk(*b)
    f.reset_tppr()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for b in batches:
        f.streaming_topk(*b)
    call_s = time.perf_counter() - t0
    call_edges = sum(len(b[2]) for b in batches)
    # the call the launcher's levels 1-3 make: rows left on the device (streaming_topk_raw), per call and on an
    # announced pass (tppr_finder.schedule_pass: the batches streamed 64 per launch ahead of the calls)
    dev_s = {}
    hi_e = min(E, nb_call * BS)
    for mode in ("per_call", "announced") * 2:          # first round untimed (kernel loading, allocator warm-up)
        f.reset_tppr()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if mode == "announced":
            f.schedule_pass(st.src[:hi_e], st.dst[:hi_e], neg[:hi_e], st.ts[:hi_e], st.eidx[:hi_e], BS, 64)
        for b in batches:
            f.streaming_topk_raw(*b)
        torch.cuda.synchronize()
        dev_s[mode] = time.perf_counter() - t0
    f.sync_check()

    t = torch.tensor([ms, e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_s = float(t[0]), float(t[1])
    if rank == 0:
        peak, peak_src = peaks()
        value = (1 if split else world) * E * args.steps / (ms * 1e-3)
        per_launch_ms = ms / max(launches, 1)
        alg_per_launch = E * bytes_per_edge() / n_launch
        achieved = alg_per_launch / (per_launch_ms * 1e-3) / 1e9
        traffic, traffic_src = (args.traffic, "--traffic") if args.traffic is not None else ncu_traffic()
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if split else "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_text(st),
                       "edges_per_gpu": E,
                       "step": "reset tables + stream the whole span once",
                       "batches_per_launch": window, "parallelism": (f"one (alpha,beta) instance per GPU x{world}, same stream, no exchange" if split else
                                                                       f"replicas x{world}") if world > 1 else "single GPU",
                       "l2": f"tables {len(al) * N * f._record_bytes / 1e6:.0f} MB + outputs exceed the 126 MB L2 (inputs larger than L2)",
                       "tie_mode": "exact (numba argsort order replayed inside groups of equal scores; bit-identical to the reference)",
                       "tie_path_prunes_last_step": tie_prunes,
                       "chain_nodes": chain_info},
            "e2e": {"value": (1 if split else world) * e2e_edges * args.steps / e2e_s, "unit": UNIT,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "step": f"all {plan.n_batches} batches through tppr_finder.stream_epoch (host numpy in, host numpy out per "
                            f"batch; windows of {args.e2e_window} batches pipelined: host dependency pass, H2D, kernel, D2H)",
                    "launches_per_step": launches_e2e,
                    "per_call_drop_in": {"value": (1 if split else world) * call_edges / call_s, "unit": UNIT,
                                         "what": f"first {nb_call} batches, one synchronous streaming_topk call per batch"},
                    "per_call_device_rows": {"value": (1 if split else world) * call_edges / dev_s["per_call"],
                                             "announced_pass": (1 if split else world) * call_edges / dev_s["announced"], "unit": UNIT,
                                             "what": "same batches through streaming_topk_raw (rows stay in HBM: what the launcher's "
                                                     "levels 1-3 call), per call and with the pass announced to the finder first"}},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak * (world if split else 1), "unit": "GB/s",
                         "frac": achieved / (peak * (world if split else 1)),
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "kernel": "tppr_span_kernel",
                         "bytes_per_edge": bytes_per_edge(), "algorithmic_bytes_per_launch": alg_per_launch,
                         "launch_ms": per_launch_ms,
                         # what actually bounds this stream: the hottest node's edges are strictly sequential
                         "critical_path": {"hottest_node_edges": hub_edges, "us_per_step": 1e3 * (ms / args.steps) / max(hub_edges, 1),
                                           "what": "time per pass / edges of the hottest node: an upper bound on the latency of one "
                                                   "step of its dependency chain (every update needs the previous one)"}},
        }
        if stable_ms is not None:
            line["tie_modes"] = {"exact": {"value": value / (1 if split else world), "ms_per_step": ms / args.steps},
                                 "stable": {"value": E / (stable_ms * 1e-3), "ms_per_step": stable_ms,
                                            "what": "equal scores keep insertion order (oracle tie_mode 1, tests/test_gpu_ties.py)"
                                                    + ("; rank 0's instance only" if split else "")}}
        if world == 1 and not args.no_flat:
            # the same kernel on a hub-free stream of the same size class (uniform endpoints): no long
            # dependency chain, i.e. the kernel's throughput regime, where HBM is the relevant roofline
            from zebra_b200 import synth
            nf = min(E, 4_000_000)
            shp = synth.StreamShape("flat", st.shape.n_nodes, 0, nf, 0, 2320.0, 1.0 / st.shape.n_nodes, 0.0)
            fs_ = synth.make_stream("flat", seed=11, shape=shp)
            fneg = synth.make_negatives(fs_, BS)
            fplan = f.make_plan(fs_.src, fs_.dst, fneg, fs_.ts, fs_.eidx, BS)
            best = 1e30
            for _ in range(3):
                f.reset_tppr()
                fplan.run(0, fplan.n_batches, window, out=out, out_slots=window)
                f.sync_check()
                best = min(best, f.last_kernel_ms())
            fplan.close()
            fa = nf * bytes_per_edge() / (best * 1e-3) / 1e9
            line["roofline"]["hub_free_stream"] = {
                "what": f"same kernel, {nf} edges with uniform endpoints over {st.shape.n_nodes} nodes (no hub chain), rows included",
                "edges_per_sec": nf / (best * 1e-3), "achieved": fa, "frac": fa / peak}
        if not args.no_cpu and world == 1:
            v, n, dt = cpu_port_throughput(st, neg, args.cpu_sample_edges, 1)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "of_cores": os.cpu_count(), "kind": "port",
                                    "sample": f"first {n} edges of the workload, bs={BS}, with negatives ({dt:.1f} s)",
                                    "reference_numba": (numba_reference_throughput(args.numba_sample_edges) if not args.no_numba
                                                        else {"unavailable": "--no-numba"})}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--edges", type=int, default=None, help="truncate the stream (debug only)")
    ap.add_argument("--workload", default="wikitalk", choices=WORKLOADS, help="shape of a BASELINE.json config (default: configs[4])")
    ap.add_argument("--batches-per-launch", type=int, default=256)
    ap.add_argument("--numba-sample-edges", type=int, default=50_000, help="prefix streamed through the reference's own numba finder")
    ap.add_argument("--no-numba", action="store_true")
    ap.add_argument("--no-stable", action="store_true", help="skip the tie_mode=stable leg")
    ap.add_argument("--shard-batches-per-launch", type=int, default=256, help="N>1: batches per launch PER GPU")
    ap.add_argument("--parity-batches", type=int, default=250, help="N>1: batches checked against the oracle before timing")
    ap.add_argument("--e2e-batches", type=int, default=2000)
    ap.add_argument("--e2e-window", type=int, default=256)
    ap.add_argument("--cpu-sample-edges", type=int, default=1_000_000)
    ap.add_argument("--ref-sample-edges", type=int, default=1_000_000)
    ap.add_argument("--traffic", type=float, default=None, help="ncu dram bytes per launch, if captured")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-flat", action="store_true", help="skip the hub-free throughput leg")
    ap.add_argument("--mode", default="sharded", choices=["sharded", "replicas", "instances"],
                    help="N>1: node-sharded state with in-kernel NVLink exchange (default), independent replicas, or one "
                         "(alpha,beta) instance per GPU (needs --gpus = number of instances)")
    ap.add_argument("--no-strong", action="store_true", help="N>1: skip the strong-scaling leg (same stream as N=1, sharded)")
    args = ap.parse_args()
    global WORKLOAD
    WORKLOAD = args.workload
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
