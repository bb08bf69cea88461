"""GPU parity of the fused neighbour gather (zebra_gather_fwd/bwd through the C ABI):
(1) against the plain torch fp32 ops of the reference (embedding_module.py:243-271) on the same inputs,
(2) end to end against recordings of the reference's GraphDiffusionEmbedding (tests/golden/embedding_*.npz).
Tolerance: fp32, rtol 1e-5 / atol 1e-6 for the gather (cosf and a 20-term sum), 2e-5 after the two
nn.Linear layers (cuBLAS vs the CPU GEMM the recording was made with)."""
import types

import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu


def _torch_reference(memory, ef, tw, tb, ids, eidx, dt, w, src):
    """The reference's ops, verbatim, on the GPU in fp32."""
    node_features = memory[ids.long()]
    edge_features = ef[eidx.long(), :]
    time_embeddings = torch.cos(dt.unsqueeze(2) * tw.view(1, 1, -1) + tb.view(1, 1, -1))
    X = torch.cat([node_features, edge_features, time_embeddings], dim=-1)
    s = torch.sum(w, dim=1)
    wn = w / s.unsqueeze(1)
    wn[s == 0] = 0
    return X, wn, memory[src.long()]


@pytest.mark.parametrize("R,k,n_tppr,Dm,De,Dt", [(600, 20, 2, 100, 172, 100), (600, 20, 2, 100, 1, 100),
                                                 (90, 7, 3, 36, 4, 20), (33, 1, 1, 10, 3, 5), (200, 31, 1, 100, 4, 100)])
def test_gather_forward_backward_vs_torch(R, k, n_tppr, Dm, De, Dt):
    from zebra_b200.embedding import FusedGather
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(R + k)
    N, E = 5000, 8000
    memory = torch.randn(N, Dm, generator=g).to(dev).requires_grad_(True)
    ef = torch.randn(E, De, generator=g).to(dev)
    tw = torch.from_numpy(1 / 10 ** np.linspace(0, 9, Dt, dtype=np.float32)).to(dev)
    tb = (torch.randn(Dt, generator=g) * 0.1).to(dev)
    ids = torch.randint(0, N, (n_tppr, R, k), generator=g, dtype=torch.int32)
    eidx = torch.randint(0, E, (n_tppr, R, k), generator=g, dtype=torch.int32)
    dt = (torch.rand(n_tppr, R, k, generator=g) * 2.6e6).float()
    w = torch.rand(n_tppr, R, k, generator=g).float()
    w[:, ::7] = 0                                   # rows whose weights sum to zero
    ids[:, ::7] = 0
    w[:, 1::5, k // 2:] = 0                          # padded slots
    src = torch.randint(0, N, (R,), generator=g, dtype=torch.int32).to(dev)
    topk = torch.cat([torch.cat([ids[j].reshape(-1).view(torch.uint8), eidx[j].reshape(-1).view(torch.uint8),
                                 dt[j].reshape(-1).view(torch.uint8), w[j].reshape(-1).view(torch.uint8)])
                      for j in range(n_tppr)]).to(dev)
    X, wn, srows = FusedGather.apply(memory, topk, src, ef, tw, tb, n_tppr, R, k)
    gX = torch.randn(X.shape, generator=g).to(dev)
    gS = torch.randn(srows.shape, generator=g).to(dev)
    # gradient reaching X is zero wherever the normalised weight is zero (it is multiplied by it downstream)
    gX = gX * (wn.reshape(n_tppr, R * k, 1) != 0)
    (X * gX).sum().add((srows * gS).sum()).backward()
    grad = memory.grad.clone()
    memory.grad = None
    # same order of creation as the reference's model: the src gather first, then the instances
    # (embedding_module.py:243-253) — autograd's accumulation order depends on it
    ref_g = (memory[src.long()] * gS).sum()
    for j in range(n_tppr):
        Xr, wr, sr = _torch_reference(memory, ef, tw, tb, ids[j].to(dev), eidx[j].to(dev), dt[j].to(dev), w[j].to(dev), src)
        assert torch.equal(X[j][:, :Dm + De], Xr.reshape(R * k, -1)[:, :Dm + De])
        assert torch.equal(X[j][:, Dm + De:], Xr.reshape(R * k, -1)[:, Dm + De:]), "time encoding must be bit-identical to torch.cos"
        torch.testing.assert_close(wn[j], wr, rtol=1e-5, atol=1e-7)
        assert torch.equal(srows, sr)
        ref_g = ref_g + (Xr.reshape(R * k, -1) * gX[j]).sum()
    ref_g.backward()
    # default (sorted) backward replays autograd's arithmetic: bit-identical gradient
    assert torch.equal(grad, memory.grad), f"{int((grad != memory.grad).sum())} gradient elements differ from autograd"
    from zebra_b200 import _lib
    assert _lib.lib().zebra_gather_check(None) == 0


def test_gather_reports_out_of_range_ids():
    from zebra_b200.embedding import FusedGather
    from zebra_b200 import _lib
    dev = torch.device("cuda")
    memory, ef = torch.zeros(10, 8, device=dev), torch.zeros(5, 4, device=dev)
    tw, tb = torch.ones(4, device=dev), torch.zeros(4, device=dev)
    ids = torch.tensor([[3, 99]], dtype=torch.int32)
    z = torch.zeros(1, 2, dtype=torch.int32)
    topk = torch.cat([ids.view(-1).view(torch.uint8), z.view(-1).view(torch.uint8), z.view(-1).view(torch.uint8),
                      z.view(-1).view(torch.uint8)]).to(dev)
    FusedGather.apply(memory, topk, torch.zeros(1, dtype=torch.int32, device=dev), ef, tw, tb, 1, 1, 2)
    assert _lib.lib().zebra_gather_check(None) != 0
    assert b"out of range" in _lib.lib().zebra_last_error()
    assert _lib.lib().zebra_gather_check(None) == 0


@pytest.mark.parametrize("strategy,name,n_edges,d_e", H.EMBED_CASES)
def test_embedding_matches_reference_recording(strategy, name, n_edges, d_e):
    from zebra_b200.embedding import GraphDiffusionEmbedding
    from zebra_b200.neighbor_finder import NeighborFinder
    g = H.load(f"embedding_{strategy}_{name}")
    st, neg, N, ef, memory = H.embedding_inputs(name, n_edges, d_e)
    assert N == int(g["num_nodes"]) and np.isclose(float(memory.sum()), float(g["mem_sum"])) and np.isclose(float(ef.sum()), float(g["ef_sum"]))
    dev = torch.device("cuda")
    args = types.SimpleNamespace(alpha_list=[0.1, 0.1], beta_list=[0.5, 0.95], topk=20, tppr_strategy=strategy,
                                 n_degree=10, n_layer=2)
    a = int(g["adj_edges"])
    nf = NeighborFinder.from_edges(st.src[:a], st.dst[:a], st.eidx[:a], st.ts[:a]) if strategy == "pruning" else None
    mod = GraphDiffusionEmbedding(torch.from_numpy(ef).to(dev), nf, d_e, 100, 100, dev, args, N).to(dev)
    sd = {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd.")}
    mod.load_state_dict(sd)
    mod.eval()
    mem = torch.from_numpy(memory).to(dev)
    with torch.no_grad():
        for b, lo in enumerate(range(0, n_edges, 200)):
            hi = lo + 200
            nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]])
            ts = np.concatenate([st.ts[lo:hi]] * 3)
            emb = mod.compute_embedding_tppr_ensemble(mem, nodes, ts, st.eidx[lo:hi].astype(np.int64), None, False)
            if f"emb_{b}" in g.files:
                ref = torch.from_numpy(g[f"emb_{b}"]).to(dev)
                assert emb.shape == ref.shape == (600, 300)
                torch.testing.assert_close(emb, ref, rtol=2e-5, atol=2e-5)


def test_gather_backward_sorted_is_reproducible_and_matches_atomic():
    """The default (sorted) backward gives bit-identical gradients run after run; the atomic variant agrees
    with it to fp32 rounding.  Hub-like ids (many slots on few nodes) make the summation order matter."""
    from zebra_b200 import embedding as E
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(5)
    R, k, n_tppr, Dm, De, Dt, N, Ee = 600, 20, 2, 100, 4, 100, 300, 500
    memory = torch.randn(N, Dm, generator=g).to(dev).requires_grad_(True)
    ef = torch.randn(Ee, De, generator=g).to(dev)
    tw = torch.from_numpy(1 / 10 ** np.linspace(0, 9, Dt, dtype=np.float32)).to(dev)
    tb = torch.zeros(Dt, device=dev)
    ids = torch.randint(0, 12, (n_tppr, R, k), generator=g, dtype=torch.int32)          # 12 hot nodes
    ids[:, ::3] = torch.randint(0, N, (n_tppr, (R + 2) // 3, k), generator=g, dtype=torch.int32)
    eidx = torch.randint(0, Ee, (n_tppr, R, k), generator=g, dtype=torch.int32)
    dt = (torch.rand(n_tppr, R, k, generator=g) * 1e5).float()
    w = torch.rand(n_tppr, R, k, generator=g).float()
    w[:, 1::5, k // 2:] = 0
    src = torch.randint(0, N, (R,), generator=g, dtype=torch.int32).to(dev)
    topk = torch.cat([torch.cat([x[j].reshape(-1).view(torch.uint8) for x in (ids, eidx, dt, w)]) for j in range(n_tppr)]).to(dev)
    gX = gS = None

    def grad(reproducible):
        nonlocal gX, gS
        old = E.REPRODUCIBLE_BACKWARD
        E.REPRODUCIBLE_BACKWARD = reproducible
        try:
            X, wn, srows = E.FusedGather.apply(memory, topk, src, ef, tw, tb, n_tppr, R, k)
            if gX is None:
                gX = torch.randn(X.shape, generator=g).to(dev) * (wn.reshape(n_tppr, R * k, 1) != 0)
                gS = torch.randn(srows.shape, generator=g).to(dev)
            memory.grad = None
            (X * gX).sum().add((srows * gS).sum()).backward()
            return memory.grad.clone()
        finally:
            E.REPRODUCIBLE_BACKWARD = old

    a, b = grad(True), grad(True)
    assert torch.equal(a, b), "sorted backward must be bit-reproducible"
    c = grad(False)
    torch.testing.assert_close(a, c, rtol=1e-4, atol=1e-3)
    # against torch's own (sort-based) backward of the same gathers
    memory.grad = None
    ref = (memory[src.long()] * gS).sum()
    for j in range(n_tppr):
        Xr, _, _ = _torch_reference(memory, ef, tw, tb, ids[j].to(dev), eidx[j].to(dev), dt[j].to(dev), w[j].to(dev), src)
        ref = ref + (Xr.reshape(R * k, -1) * gX[j]).sum()
    ref.backward()
    assert torch.equal(a, memory.grad), "sorted backward must equal torch autograd bit for bit"
