"""Raw-message construction of a batch on the device.

`get_raw_messages` has the signature and return value of the reference's `TGN.get_raw_messages`
(model/tgn_model.py:204-226): the last interaction of every distinct node of [src ; dst] is selected on the
host exactly as the reference does it (np.flip + np.unique(return_index) on 2n ids — the host-side flag array
`Memory.nodes` needs those ids on the host anyway), then ONE pinned host-to-device copy carries (node, partner,
edge id, float32 time) and ONE kernel (zebra_raw_messages) writes
[memory[u] | memory[v] | edge_features[e] | cos((t - last_update[u]) w + b)] — instead of four index
operations with their own copies, a time encoding and a torch.cat.  `install()` (launcher level 3) patches
the reference's TGN class.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ._lib import check, lib


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def raw_messages(idx_dev, n_rows, memory, last_update, edge_features, time_w, time_b):
    """idx_dev: int32 CUDA tensor [4, n_rows] = node, partner, edge id, float32 time bits."""
    Dm, De, Dt = memory.shape[1], edge_features.shape[1], time_w.numel()
    out = torch.empty((n_rows, 2 * Dm + De + Dt), dtype=torch.float32, device=memory.device)
    if n_rows:
        with torch.cuda.device(memory.device):
            check(lib().zebra_raw_messages(_ptr(idx_dev[0]), _ptr(idx_dev[1]), _ptr(idx_dev[2]), _ptr(idx_dev[3]), n_rows,
                                           _ptr(memory), memory.shape[0], Dm, _ptr(last_update), _ptr(edge_features),
                                           edge_features.shape[0], De, _ptr(time_w), _ptr(time_b), Dt, _ptr(out),
                                           C.c_void_p(torch.cuda.current_stream(memory.device).cuda_stream)))
    return out


def get_raw_messages(self, source_nodes, source_node_embedding, destination_nodes, destination_node_embedding, edge_times, edge_idxs):
    if self.use_source_embedding_in_message or self.use_destination_embedding_in_message or not self.memory.memory.is_cuda:
        return _REF["get_raw_messages"](self, source_nodes, source_node_embedding, destination_nodes, destination_node_embedding,
                                        edge_times, edge_idxs)
    # model/tgn_model.py:206-211, unchanged: last occurrence of every node, ascending node order
    reversed_source_nodes = np.flip(source_nodes)
    unique_source_nodes, reversed_index = np.unique(reversed_source_nodes, return_index=True)
    index = len(source_nodes) - reversed_index - 1
    U = len(unique_source_nodes)
    stage = self.__dict__.get("_zebra_msg_stage")
    if stage is None or stage.shape[1] < U:
        stage = self.__dict__["_zebra_msg_stage"] = torch.empty((4, max(U, 1024)), dtype=torch.int32).pin_memory()
    h = stage.numpy()
    h[0, :U] = unique_source_nodes
    h[1, :U] = destination_nodes[index]
    h[2, :U] = edge_idxs[index]
    h[3, :U].view(np.float32)[:] = edge_times[index]          # float64 -> float32 on the host, as torch.from_numpy(...).float()
    mem = self.memory
    dev = mem.memory.device
    idx = stage[:, :U].to(dev, non_blocking=True)             # the previous batch's copy is long done (loss.item() synchronises)
    te = self.time_encoder
    msg = raw_messages(idx, U, mem.memory, mem.last_update, self.edge_raw_features, te.w.weight.detach().reshape(-1).contiguous(),
                       te.w.bias.detach().contiguous())
    return unique_source_nodes, msg, idx[3].view(torch.float32)


_REF = {}


def install():
    import model.tgn_model as tm                  # the reference's module (reference root is on sys.path)
    if "get_raw_messages" not in _REF:
        _REF["get_raw_messages"] = tm.TGN.get_raw_messages
    tm.TGN.get_raw_messages = get_raw_messages
