"""Golden fixture for the neighbour gather: EXECUTES THE REFERENCE's GraphDiffusionEmbedding
(/root/reference/modules/embedding_module.py:76-278, model/time_encoding.py) on the CPU, unmodified,
with seeded frozen weights in eval mode, for both --tppr_strategy values, and records the inputs, the
module's state_dict and the returned embeddings to tests/golden/embedding_*.npz.

Run (build container only):  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_embedding.py
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

from zebra_b200 import synth  # noqa: E402


def main():
    import torch
    sys.path.insert(0, "/root/reference")
    from modules.embedding_module import GraphDiffusionEmbedding
    from model.time_encoding import TimeEncode
    from utils.util import get_neighbor_finder

    from tests import helpers as H
    for strategy, name, n_edges, d_e in H.EMBED_CASES:
        torch.manual_seed(5)
        st, neg, N, ef, memory = H.embedding_inputs(name, n_edges, d_e)
        args = types.SimpleNamespace(alpha_list=[0.1, 0.1], beta_list=[0.5, 0.95], topk=20, tppr_strategy=strategy,
                                     n_degree=10, n_layer=2)

        class D:
            sources, destinations, edge_idxs, timestamps = st.src[:1800], st.dst[:1800], st.eidx[:1800], st.ts[:1800]
        nf = get_neighbor_finder(D) if strategy == "pruning" else None
        mod = GraphDiffusionEmbedding(node_features=None, edge_features=torch.from_numpy(ef), memory=None,
                                      neighbor_finder=nf, time_encoder=TimeEncode(100), n_layers=2,
                                      n_node_features=100, n_edge_features=d_e, n_time_features=100,
                                      embedding_dimension=100, device="cpu", n_heads=2, dropout=0.1, use_memory=True,
                                      args=args, num_nodes=N)
        mod.eval()
        mem_t = torch.from_numpy(memory)
        outs = {}
        batches = [(lo, lo + 200) for lo in range(0, n_edges, 200)]
        keep = {len(batches) - 1, len(batches) - 2, 3}
        with torch.no_grad():
            for b, (lo, hi) in enumerate(batches):
                nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]])
                ts = np.concatenate([st.ts[lo:hi]] * 3)
                emb = mod.compute_embedding_tppr_ensemble(mem_t, nodes, ts, st.eidx[lo:hi].astype(np.int64), None, False)
                if b in keep:
                    outs[f"emb_{b}"] = emb.numpy().copy()
        sd = {f"sd.{k}": v.numpy() for k, v in mod.state_dict().items()}
        path = os.path.join(HERE, f"embedding_{strategy}_{name}.npz")
        np.savez_compressed(path, num_nodes=N, strategy=strategy, adj_edges=1800, mem_sum=float(memory.sum()),
                            ef_sum=float(ef.sum()), **sd, **outs)
        print(path, os.path.getsize(path) // 1024, "KiB", list(outs))


if __name__ == "__main__":
    main()
