"""Small, fixed workloads for `ncu` captures of the span kernel (tools/gpu_check.sh ncu):
    python tools/ncu_targets.py hub    # Wiki-Talk-shaped prefix: bound by the hottest node's chain
    python tools/ncu_targets.py flat   # uniform endpoints: the kernel's throughput regime
Each streams a resident plan in launches of 256 batches, twice."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from zebra_b200 import synth
from zebra_b200.finder import tppr_finder

mode = sys.argv[1] if len(sys.argv) > 1 else "hub"
n = 512_000
if mode == "hub":
    st = synth.make_stream("wikitalk", n_edges=n)
else:
    shp = synth.StreamShape("flat", 1_140_149, 0, n, 0, 2320.0, 1.0 / 1_140_149, 0.0)
    st = synth.make_stream("flat", seed=11, shape=shp)
neg = synth.make_negatives(st, 200)
f = tppr_finder(st.shape.table_nodes, 20, 2, [0.1, 0.1], [0.5, 0.95])
plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, 200)
out = plan.alloc_out(256)
for rep in range(2):
    f.reset_tppr()
    plan.run(0, plan.n_batches, 256, out=out, out_slots=256)
    f.sync_check()
    print(mode, rep, f"{f.last_kernel_ms():.2f} ms", f.counters(), flush=True)
