"""Per-batch cost of the T-PPR query inside a validation-like pass: one synchronous-free `streaming_topk_raw`
call per batch (the launcher's level-1..3 path) against the same calls on an announced pass
(`tppr_finder.schedule_pass`).  Prints one JSON line.

    python tools/perf_schedule.py [--workload reddit] [--edges 120000] [--window 64]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from zebra_b200 import synth
    from zebra_b200.finder import tppr_finder
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="reddit")
    ap.add_argument("--edges", type=int, default=120_000)
    ap.add_argument("--window", type=int, default=64)
    a = ap.parse_args()
    st = synth.make_stream(a.workload, n_edges=a.edges)
    bs = 200
    neg = synth.make_negatives(st, bs)
    N = st.shape.table_nodes
    split = int(0.7 * a.edges) // bs * bs
    nb = (a.edges - split + bs - 1) // bs
    calls = []
    for b in range(nb):
        lo, hi = split + b * bs, min(a.edges, split + (b + 1) * bs)
        calls.append((np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32),
                      np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi]))
    res = {}
    digest = {}
    for mode in ("per_call", "announced") * 3:
        f = tppr_finder(N, 20, 2, [0.1, 0.1], [0.5, 0.95])
        f.compute_val_tppr(st.src[:split], st.dst[:split], st.ts[:split], st.eidx[:split])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if mode == "announced":
            f.schedule_pass(st.src[split:], st.dst[split:], neg[split:], st.ts[split:], st.eidx[split:], bs, a.window)
        acc = torch.zeros((), dtype=torch.float64, device="cuda")
        for c in calls:
            buf, rows = f.streaming_topk_raw(*c)
            acc += buf[rows * 20 * 12: rows * 20 * 16].view(torch.float32).sum()     # a consumer on the stream
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        res[mode] = min(res.get(mode, 1e9), dt)
        digest[mode] = float(acc)
        f.sync_check()
    assert digest["per_call"] == digest["announced"], digest
    print(json.dumps({"workload": f"{a.workload}-shaped, {a.edges} edges, pass = last {a.edges - split} edges in {nb} batches of {bs}",
                      "per_call_us_per_batch": 1e6 * res["per_call"] / nb, "announced_us_per_batch": 1e6 * res["announced"] / nb,
                      "window": a.window, "speedup": res["per_call"] / res["announced"],
                      "per_call_edges_per_s": (a.edges - split) / res["per_call"],
                      "announced_edges_per_s": (a.edges - split) / res["announced"]}))


if __name__ == "__main__":
    main()
