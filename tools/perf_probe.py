"""Quick timing probe (not the bench): per-batch streaming and bulk update-only throughput."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from zebra_b200 import synth
from zebra_b200.finder import tppr_finder

def per_batch(name, n_edges, bs=200):
    st = synth.make_stream(name, n_edges=n_edges)
    neg = synth.make_negatives(st, bs)
    f = tppr_finder(st.shape.table_nodes, 20, 2, [0.1, 0.1], [0.5, 0.95])
    batches = []
    for lo in range(0, n_edges, bs):
        hi = min(lo + bs, n_edges)
        batches.append((np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32),
                        np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi]))
    for mode in ("host", "device"):
        f.reset_tppr()
        for b in batches[:20]:
            f.streaming_topk(*b)
        f.reset_tppr()
        torch.cuda.synchronize(); t = time.perf_counter(); kms = 0.0
        for b in batches:
            if mode == "host":
                f.streaming_topk(*b); kms += f.last_kernel_ms()
            else:
                f.streaming_topk_device(*b)
        f.sync_check()
        dt = time.perf_counter() - t
        print(f"{name} bs={bs} {mode}: {n_edges/dt/1e6:.3f} M edges/s  ({dt/len(batches)*1e6:.1f} us/batch, kernel {kms/len(batches)*1e3:.1f} us/batch) ties={f.counters()}", flush=True)

def bulk(name, n_edges=None):
    st = synth.make_stream(name, n_edges=n_edges)
    E = st.n_edges
    f = tppr_finder(st.shape.table_nodes, 20, 2, [0.1, 0.1], [0.5, 0.95])
    src, dst = st.src.astype(np.int64), st.dst.astype(np.int64)
    f.compute_val_tppr(src[:100000], dst[:100000], st.ts[:100000], st.eidx[:100000])
    for _ in range(2):
        f.reset_tppr()
        torch.cuda.synchronize(); t = time.perf_counter()
        f.compute_val_tppr(src, dst, st.ts, st.eidx)
        dt = time.perf_counter() - t
        print(f"{name} bulk update-only E={E}: {E/dt/1e6:.2f} M edges/s wall, kernel-span {f.last_kernel_ms():.1f} ms -> {E/f.last_kernel_ms()/1e3:.2f} M edges/s, GB/s alg={E*2*1976/f.last_kernel_ms()/1e6:.1f} {f.counters()}", flush=True)

def flat(n_nodes=1_140_149, n_edges=4_000_000, bs=200, window=64):
    """No hubs (uniform endpoints): no dependency chain longer than a few edges, so this is the span
    kernel's THROUGHPUT, the regime where the HBM roofline is the relevant bound."""
    shape = synth.StreamShape("flat", n_nodes, 0, n_edges, 0, 2320.0, 1.0 / n_nodes, 0.0)
    st = synth.make_stream("flat", seed=11, shape=shape)
    neg = synth.make_negatives(st, bs)
    f = tppr_finder(st.shape.table_nodes, 20, 2, [0.1, 0.1], [0.5, 0.95])
    plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, bs)
    out = plan.alloc_out(window)
    for rep in range(3):
        f.reset_tppr()
        plan.run(0, plan.n_batches, window, out=out, out_slots=window)
        f.sync_check()
        ms = f.last_kernel_ms()
        print(f"flat stream E={n_edges} N={n_nodes} with rows: {n_edges/ms/1e3:.1f} M edges/s, alg {n_edges*6856/ms/1e6:.0f} GB/s "
              f"({n_edges*6856/ms/1e6/6534.1*100:.1f}% of measured HBM peak) {f.counters()}", flush=True)


def star(n_edges=100000, n_nodes=200000, pool=None, insts=((1, [0.1], [0.5]), (1, [0.1], [0.95]), (2, [0.1, 0.1], [0.5, 0.95]))):
    """Every edge touches node 1: throughput = 1 / (latency of one step of a hub chain)."""
    rng = np.random.default_rng(0)
    src = np.ones(n_edges, np.int64)
    dst = rng.integers(2, pool or n_nodes, n_edges).astype(np.int64)
    ts = np.arange(n_edges, dtype=np.float64)
    eidx = np.arange(1, n_edges + 1, dtype=np.int64)
    for n_t, al, be in insts:
        f = tppr_finder(n_nodes, 20, n_t, al, be)
        f.compute_val_tppr(src[:1000], dst[:1000], ts[:1000], eidx[:1000])
        f.reset_tppr()
        f.compute_val_tppr(src, dst, ts, eidx)
        ms = f.last_kernel_ms()
        pr = f.profile()
        if pr.any():
            names = ["claim+load", "wait", "upd_hot", "upd_2nd", "emit", "c.get+push", "c.prefetch", "c.update", "u.dup", "u.pos", "u.rank", "u.tie", "u.store"]
            print("   cycles/item:", {nm: round(float(v) / (n_edges * n_t), 1) for nm, v in zip(names, pr) if nm})
        print(f"star chain pool={pool} n_tppr={n_t} beta={be}: {ms*1e3/n_edges:.2f} us per hub step ({n_edges/ms/1e3:.2f} M edges/s) {f.counters()}", flush=True)


if __name__ == "__main__":
    star(insts=((1, [0.1], [0.5]), (1, [0.1], [0.95])))
    star(pool=2000, insts=((1, [0.1], [0.5]), (1, [0.1], [0.95])))
    if os.environ.get("ZEBRA_TPPR_LIB") or "--star" in sys.argv:
        sys.exit(0)
    if "--flat" in sys.argv:
        flat()
        sys.exit(0)
    per_batch("wikipedia", 157474)
    per_batch("reddit", 200000)
    per_batch("wikitalk", 400000)
    bulk("reddit")
    bulk("wikitalk")
    flat()
