"""Golden fixture for the TRAIN branch of the embedding (lazy memory update + gather + backward): EXECUTES THE
REFERENCE's GraphDiffusionEmbedding.compute_embedding_tppr_ensemble(train=True), its Memory and its
GRUMemoryUpdater.get_updated_memory (/root/reference/modules/embedding_module.py:217-278,
modules/memory.py, modules/memory_updater.py:61-90) on the CPU, unmodified, with seeded weights and dropout
switched off, and records the embeddings and a few parameter gradients of selected batches.

Run (build container only):  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_embedding_train.py
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

GRAD_KEYS = ("mod.fc1.weight", "mod.fc2.bias", "gru.memory_updater.bias_ih", "gru.memory_updater.bias_hh")


def main():
    import torch
    sys.path.insert(0, "/root/reference")
    from modules.embedding_module import GraphDiffusionEmbedding
    from modules.memory import Memory
    from modules.memory_updater import GRUMemoryUpdater
    from model.time_encoding import TimeEncode

    from tests import helpers as H
    strategy, name, n_edges, d_e = "streaming", "wikipedia", 2400, 172
    torch.manual_seed(11)
    st, neg, N, ef, memory = H.embedding_inputs(name, n_edges, d_e)
    args = types.SimpleNamespace(alpha_list=[0.1, 0.1], beta_list=[0.5, 0.95], topk=20, tppr_strategy=strategy,
                                 n_degree=10, n_layer=2)
    mod = GraphDiffusionEmbedding(node_features=None, edge_features=torch.from_numpy(ef), memory=None,
                                  neighbor_finder=None, time_encoder=TimeEncode(100), n_layers=2,
                                  n_node_features=100, n_edge_features=d_e, n_time_features=100,
                                  embedding_dimension=100, device="cpu", n_heads=2, dropout=0.1, use_memory=True,
                                  args=args, num_nodes=N)
    mod.train()
    for layer in mod.modules():
        if isinstance(layer, torch.nn.Dropout):
            layer.p = 0.0
    d_msg = 2 * 100 + d_e + 100
    mem = Memory(n_nodes=N, memory_dimension=100, input_dimension=d_msg, message_dimension=d_msg, device="cpu")
    mem.memory = torch.from_numpy(memory.copy())
    rng = np.random.default_rng(23)
    mem.messages = torch.from_numpy((rng.standard_normal((N, d_msg)) * 0.3).astype(np.float32))
    upd = GRUMemoryUpdater(message_dimension=d_msg, memory_dimension=100, device="cpu")
    params = {**{f"mod.{k}": v for k, v in mod.named_parameters()}, **{f"gru.{k}": v for k, v in upd.named_parameters()}}
    batches = [(lo, lo + 200) for lo in range(0, n_edges, 200)]
    keep = {2, len(batches) - 1}
    flags = np.zeros((len(batches), N), dtype=bool)
    outs = {}
    for b, (lo, hi) in enumerate(batches):
        flags[b] = rng.random(N) < (0.0 if b == 0 else 0.5)
        mem.nodes = flags[b].copy()
        nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]])
        ts = np.concatenate([st.ts[lo:hi]] * 3)
        for p in params.values():
            p.grad = None
        emb = mod.compute_embedding_tppr_ensemble(mem, nodes, ts, st.eidx[lo:hi].astype(np.int64), upd, True)
        emb.square().sum().backward()
        if b in keep:
            outs[f"emb_{b}"] = emb.detach().numpy().copy()
            for k in GRAD_KEYS:
                outs[f"grad_{b}.{k}"] = params[k].grad.numpy().copy()
    sd = {f"sd.{k}": v.numpy() for k, v in mod.state_dict().items()}
    sd.update({f"gru_sd.{k}": v.numpy() for k, v in upd.memory_updater.state_dict().items()})
    path = os.path.join(HERE, f"embedding_train_{strategy}_{name}.npz")
    np.savez_compressed(path, num_nodes=N, n_edges=n_edges, d_e=d_e, flags=np.packbits(flags, axis=1), msg_seed=23,
                        mem_sum=float(memory.sum()), **sd, **outs)
    print(path, os.path.getsize(path) // 1024, "KiB", sorted(outs)[:4], "...")


if __name__ == "__main__":
    main()
