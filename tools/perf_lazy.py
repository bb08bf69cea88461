"""Per-batch cost of the train-branch memory update + neighbour gather (forward and backward), level 1
(reference's get_updated_memory: D2H of the ids, host unique, clone of the whole memory, index_put; then the
fused gather) against level 2 (zebra_pending_select + updater on the pending rows + zebra_gather_*_lazy).
Wall clock around a synchronised batch, because the host work is part of what level 2 removes.

    python tools/perf_lazy.py
"""
import os
import sys
import time
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from zebra_b200 import embedding as E


def get_updated_memory(cell, memory, index):
    """modules/memory_updater.py:61-90 restated (the part whose result the embedding uses, plus the
    last_update clone the reference also pays for)."""
    flags = memory.nodes
    ids = index[np.where(flags[index])[0]]
    if len(ids) == 0:
        return memory.memory.clone(), memory.last_update.clone()
    updated = memory.memory.clone()
    updated[ids] = cell(memory.messages[ids], updated[ids])
    last = memory.last_update.clone()
    last[ids] = memory.timestamps[ids]
    return updated, last


def run(N, De, tag, R=600, k=20, n_tppr=2, Dm=100, Dt=100, batches=30, hot=4000):
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(1)
    Dmsg = 2 * Dm + De + Dt
    mem = types.SimpleNamespace(memory=torch.randn(N, Dm, device=dev), messages=torch.randn(N, Dmsg, device=dev),
                                last_update=torch.zeros(N, device=dev), timestamps=torch.zeros(N, device=dev),
                                nodes=np.zeros(N, dtype=bool))
    Ee = 100000
    ef = torch.randn(Ee, De, device=dev)
    tw = torch.from_numpy(1 / 10 ** np.linspace(0, 9, Dt, dtype=np.float32)).to(dev)
    tb = torch.zeros(Dt, device=dev)
    cell = torch.nn.GRUCell(Dmsg, Dm).to(dev)
    rng = np.random.default_rng(2)
    active = rng.choice(N, hot, replace=False)                      # recently active nodes: neighbours and flags
    holder = types.SimpleNamespace()
    res = {}
    for level in (1, 2, 1, 2):
        ts = []
        for b in range(batches):
            ids = torch.from_numpy(rng.choice(active, (n_tppr, R, k)).astype(np.int32))
            eidx = torch.randint(1, Ee, (n_tppr, R, k), generator=g, dtype=torch.int32)
            dt = (torch.rand(n_tppr, R, k, generator=g) * 2.6e6).float()
            w = torch.rand(n_tppr, R, k, generator=g).float()
            topk = torch.cat([torch.cat([x[j].reshape(-1).view(torch.uint8) for x in (ids, eidx, dt, w)]) for j in range(n_tppr)]).to(dev)
            src = torch.from_numpy(rng.choice(active, R).astype(np.int32)).to(dev)
            mem.nodes[:] = False
            mem.nodes[rng.choice(active, 400, replace=False)] = True   # what store_raw_messages leaves behind per batch
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            plane = R * k
            if level == 1:
                idp = torch.stack([topk[j * plane * 16: j * plane * 16 + plane * 4].view(torch.int32) for j in range(n_tppr)])
                index = np.unique(idp.cpu().numpy())
                memory, _ = get_updated_memory(cell, mem, index)
                X, wn, srows = E.FusedGather.apply(memory, topk, src, ef, tw, tb, n_tppr, R, k)
            else:
                idx, bitmap, wp = E.pending_select(topk, n_tppr, R, k, E._flags_to_device(holder, mem, dev), N)
                if idx.numel():
                    upd = cell(mem.messages[idx], mem.memory[idx])
                    X, wn, srows = E.FusedGatherLazy.apply(mem.memory, upd, bitmap, wp, topk, src, ef, tw, tb, n_tppr, R, k)
                else:
                    X, wn, srows = E.FusedGather.apply(mem.memory, topk, src, ef, tw, tb, n_tppr, R, k)
            loss = X.sum() + srows.sum()
            if loss.requires_grad:
                loss.backward()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
            cell.zero_grad(set_to_none=True)
        res[level] = float(np.median(ts[5:])) * 1e3
    print(f"{tag}: N={N} De={De} R={R} k={k} n_tppr={n_tppr}: level 1 {res[1]:.3f} ms/batch, level 2 {res[2]:.3f} ms/batch "
          f"({res[1] / res[2]:.2f}x), peak memory {torch.cuda.max_memory_allocated() / 1e9:.2f} GB", flush=True)


if __name__ == "__main__":
    run(10985, 172, "C2 shape (Reddit)")
    run(194086, 1, "C4 shape (SuperUser)")
    run(1140150, 1, "C5 shape (Wiki-Talk)")
