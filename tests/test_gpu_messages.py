"""zebra_raw_messages / zebra_b200.messages.get_raw_messages against the reference's own torch operations
(model/tgn_model.py:204-226, restated below line by line), bit for bit: selection of the last interaction per
node, the four gathers, the time encoding of (float32(t) - last_update[u]) and the concatenation."""
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference(self, source_nodes, destination_nodes, edge_times, edge_idxs):
    reversed_source_nodes = np.flip(source_nodes)
    unique_source_nodes, reversed_index = np.unique(reversed_source_nodes, return_index=True)
    index = len(source_nodes) - reversed_index - 1
    unique_destination_nodes = destination_nodes[index]
    edge_times = edge_times[index]
    edge_idxs = edge_idxs[index]
    edge_times = torch.from_numpy(edge_times).float().to(self.device)
    edge_features = self.edge_raw_features[edge_idxs]
    source_memory = self.memory.memory[unique_source_nodes, :]
    destination_memory = self.memory.memory[unique_destination_nodes, :]
    source_time_delta = edge_times - self.memory.last_update[unique_source_nodes,]
    enc = self.time_encoder(source_time_delta.unsqueeze(dim=1)).view(len(unique_source_nodes,), -1)
    return unique_source_nodes, torch.cat([source_memory, destination_memory, edge_features, enc], dim=1), edge_times


@pytest.mark.parametrize("N,E,De,n", [(3000, 5000, 172, 200), (500, 900, 4, 200), (40, 300, 1, 137)])
def test_raw_messages_bit_identical(N, E, De, n):
    from zebra_b200 import messages as M
    from zebra_b200.embedding import TimeEncode
    dev = torch.device("cuda")
    rng = np.random.default_rng(5)
    self = types.SimpleNamespace(
        device=dev, use_source_embedding_in_message=False, use_destination_embedding_in_message=False,
        edge_raw_features=torch.from_numpy(rng.standard_normal((E + 1, De)).astype(np.float32)).to(dev),
        time_encoder=TimeEncode(100).to(dev),
        memory=types.SimpleNamespace(memory=torch.from_numpy(rng.standard_normal((N, 100)).astype(np.float32)).to(dev),
                                     last_update=torch.from_numpy(rng.uniform(0, 2e6, N).astype(np.float32)).to(dev)))
    for _ in range(5):
        src = rng.integers(1, N, n); dst = rng.integers(1, N, n)
        t = np.sort(rng.uniform(2e6, 2.6e6, n)); e = rng.integers(1, E + 1, n)
        s2, d2 = np.concatenate([src, dst]), np.concatenate([dst, src])
        t2, e2 = np.concatenate([t, t]), np.concatenate([e, e])
        ru, rm, rt = _reference(self, s2, d2, t2, e2)
        gu, gm, gt = M.get_raw_messages(self, s2, None, d2, None, t2, e2)
        assert np.array_equal(ru, gu)
        assert torch.equal(rm, gm), "messages must be bit-identical"
        assert torch.equal(rt, gt)
