"""Developer probe: per-phase cycle counters of the chain warps on a real-shaped stream (needs the
-DZEBRA_PROFILE library: ZEBRA_TPPR_LIB=zebra_b200/lib/libzebra_tppr_prof.so)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from zebra_b200 import synth
from zebra_b200.finder import tppr_finder

name = sys.argv[1] if len(sys.argv) > 1 else "wikitalk"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000
window = int(sys.argv[3]) if len(sys.argv) > 3 else 64
st = synth.make_stream(name, n_edges=n)
neg = synth.make_negatives(st, 200)
for betas in ([0.5], [0.95], [0.5, 0.95]):
    f = tppr_finder(st.shape.table_nodes, 20, len(betas), [0.1] * len(betas), betas)
    plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, 200)
    out = plan.alloc_out(window)
    for rep in range(2):
        f.reset_tppr()
        f.profile()
        plan.run(0, plan.n_batches, window, out=out, out_slots=window)
        f.sync_check()
        ms = f.last_kernel_ms()
    c = f.counters()
    pr = f.profile()
    names = ["claim+load", "wait", "upd_hot", "upd_2nd", "emit", "c.get+push", "c.prefetch", "c.update", "u.dup", "u.pos", "u.rank", "u.tie", "u.store"]
    deg = np.bincount(np.concatenate([st.src, st.dst]))
    steps = max(int(deg.max()) * len(betas), 1)            # chain marks are taken on the hottest chain only
    items = c["items"]
    print(f"{name} E={n} betas={betas} window={window}: {ms:.1f} ms = {n/ms/1e3:.2f} M edges/s  {c}")
    if pr.any():
        print(f"   hottest chain ({int(deg.max())} steps), per step:", {nm: round(float(v) / steps, 1) for nm, v in zip(names, pr) if nm.startswith("c.")})
        print("   chain partner waits: steps that had to wait %.3f, polls per step %.2f" % (float(pr[14]) / steps, float(pr[13]) / steps))
        print("   per item      :", {nm: round(float(v) / items, 1) for nm, v in zip(names, pr) if not nm.startswith("c.")})
