"""Timing probe of the two stateless kernels next to the span kernel: the fused neighbour gather
(gather_fwd8_kernel / gather_bwd_sorted_kernel) and the pruning query (prune_topk_kernel).  CUDA events,
L2 flushed between timed launches, algorithmic bytes per SURVEY.md section 8(d).

    python tools/perf_kernels.py            # prints one line per case
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from zebra_b200 import synth
from zebra_b200.embedding import FusedGather
from zebra_b200.neighbor_finder import NeighborFinder

PEAK = 6534.1
if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")):
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])


def timed(fn, reps=20, flush=None):
    fn(); fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        if flush is not None:
            flush.add_(1)                        # 512 MB read+write: evicts the 126 MB L2
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return float(np.median(ms))


def gather(R, k=20, n_tppr=2, Dm=100, De=172, Dt=100, N=10985, E=672448, flush=None, tag=""):
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(1)
    memory = torch.randn(N, Dm, generator=g).to(dev).requires_grad_(True)
    ef = torch.randn(E, De, generator=g).to(dev)
    tw = torch.from_numpy(1 / 10 ** np.linspace(0, 9, Dt, dtype=np.float32)).to(dev)
    tb = torch.zeros(Dt, device=dev)
    ids = torch.randint(1, N, (n_tppr, R, k), generator=g, dtype=torch.int32)
    eidx = torch.randint(1, E, (n_tppr, R, k), generator=g, dtype=torch.int32)
    dt = (torch.rand(n_tppr, R, k, generator=g) * 2.6e6).float()
    w = torch.rand(n_tppr, R, k, generator=g).float()
    src = torch.randint(1, N, (R,), generator=g, dtype=torch.int32).to(dev)
    topk = torch.cat([torch.cat([x[j].reshape(-1).view(torch.uint8) for x in (ids, eidx, dt, w)]) for j in range(n_tppr)]).to(dev)
    slots = n_tppr * R * k
    fwd_bytes = slots * (4 * Dm + 4 * De + 16 + 4 * (Dm + De + Dt)) + R * 8 * Dm
    with torch.no_grad():
        ms = timed(lambda: FusedGather.apply(memory, topk, src, ef, tw, tb, n_tppr, R, k), flush=flush)
    print(f"gather_fwd{tag} R={R} k={k} n_tppr={n_tppr} De={De}: {ms*1e3:.1f} us, alg {fwd_bytes/1e6:.1f} MB -> "
          f"{fwd_bytes/ms/1e6:.0f} GB/s ({fwd_bytes/ms/1e6/PEAK*100:.1f}% of {PEAK:.0f})", flush=True)
    X, wn, srows = FusedGather.apply(memory, topk, src, ef, tw, tb, n_tppr, R, k)
    gX = torch.randn_like(X)
    gS = torch.randn_like(srows)

    def bwd():
        memory.grad = None
        torch.autograd.backward([X, srows], [gX, gS], retain_graph=True)
    ms = timed(bwd, flush=flush)
    bwd_bytes = slots * (4 * (Dm + De + Dt) // 1 * 0 + 4 * Dm * 2 + 8) + N * Dm * 4
    print(f"gather_bwd{tag} (default: stable sort of the destination ids + autograd-order sums) R={R}: {ms*1e3:.1f} us, alg {bwd_bytes/1e6:.1f} MB -> "
          f"{bwd_bytes/ms/1e6:.0f} GB/s", flush=True)


def prune(nq, name="mooc", width=10, depth=2, k=20, flush=None):
    st = synth.make_stream(name)
    nf = NeighborFinder.from_edges(st.src, st.dst, st.eidx, st.ts, num_nodes=st.shape.table_nodes)
    rng = np.random.default_rng(2)
    pick = rng.integers(st.n_edges // 2, st.n_edges, nq)
    q = st.src[pick].astype(np.int32)
    t = st.ts[pick]
    ms = timed(lambda: nf.get_pruned_topk_device(q, t, width, depth, 0.1, 0.95, k), reps=10, flush=flush)
    visited = width + width * width
    alg = nq * (16 * visited + 16 * k)
    print(f"prune_topk {name} nq={nq} width={width} depth={depth}: {ms*1e3:.1f} us incl. H2D of the queries, "
          f"{nq/ms/1e3:.2f} M queries/s, alg <= {alg/1e6:.1f} MB -> {alg/ms/1e6:.0f} GB/s", flush=True)


if __name__ == "__main__":
    flush = torch.zeros(128 * 1024 * 1024, dtype=torch.float32, device="cuda")
    gather(600, flush=flush, tag=" [one batch, C1/C2 shape]")
    gather(600, flush=None, tag=" [one batch, warm L2]")
    gather(60000, flush=flush, tag=" [100 batches]")
    gather(60000, De=1, N=1140150, E=8, flush=flush, tag=" [C5 shape, no edge feats]")
    prune(600, flush=flush)
    prune(60000, flush=flush)
