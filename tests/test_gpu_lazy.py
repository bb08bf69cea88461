"""GPU parity of the lazy memory update on the device (zebra_pending_select, zebra_gather_fwd_lazy,
zebra_gather_bwd_lazy; level 2 of zebra_b200.launch) against the reference's own sequence of torch ops
(modules/embedding_module.py:227-230 -> modules/memory_updater.py:61-90: host-side unique + flag lookup,
clone of the whole memory, index_put of the updater's rows, gathers from the clone).  Integer outputs and
the forward are compared bit for bit; so is the gradient that reaches the updater's output rows."""
import types

import numpy as np
import pytest
import torch

from zebra_b200 import synth

pytestmark = pytest.mark.gpu


def _topk(ids, eidx, dt, w, dev):
    n_tppr = ids.shape[0]
    return torch.cat([torch.cat([x[j].reshape(-1).view(torch.uint8) for x in (ids, eidx, dt, w)]) for j in range(n_tppr)]).to(dev)


@pytest.mark.parametrize("N,R,k,n_tppr,p_flag", [(33, 7, 3, 1, 0.5), (5000, 600, 20, 2, 0.3), (5000, 600, 20, 2, 0.0),
                                                 (4097, 90, 31, 3, 1.0), (1140150, 600, 20, 2, 0.4)])
def test_pending_select_matches_host_unique(N, R, k, n_tppr, p_flag):
    from zebra_b200.embedding import pending_select
    dev = torch.device("cuda")
    rng = np.random.default_rng(N + R)
    ids = rng.integers(0, N, (n_tppr, R, k)).astype(np.int32)
    ids[:, ::5, k // 2:] = 0                                      # pad slots
    ids[0, 0, 0] = N - 1                                          # last bit of the last word
    flags = rng.random(N) < p_flag
    z = torch.zeros(n_tppr, R, k, dtype=torch.int32)
    topk = _topk(torch.from_numpy(ids), z, z.view(torch.float32), z.view(torch.float32), dev)
    idx, bitmap, wordprefix = pending_select(topk, n_tppr, R, k, torch.from_numpy(flags.view(np.uint8)).to(dev), N)
    # memory_updater.py:66-70 on the index embedding_module.py:229 builds
    index = np.unique(ids)
    want = index[np.where(flags[index])[0]]
    assert idx.dtype == torch.int64 and np.array_equal(idx.cpu().numpy(), want)
    bits = np.unpackbits(bitmap.cpu().numpy().view(np.uint8), bitorder="little")[:N].astype(bool)
    assert np.array_equal(np.where(bits)[0], want)
    words = bitmap.cpu().numpy().view(np.uint32)
    pc = np.array([bin(int(x)).count("1") for x in words]) if len(words) < 5000 else \
        np.unpackbits(words.view(np.uint8)).reshape(len(words), 32).sum(1)
    assert np.array_equal(wordprefix.cpu().numpy(), np.concatenate([[0], np.cumsum(pc)[:-1]]))


@pytest.mark.parametrize("N,R,k,n_tppr,Dm,De,Dt,hot", [(5000, 600, 20, 2, 100, 172, 100, 0), (300, 600, 20, 2, 100, 4, 100, 12),
                                                        (700, 90, 7, 3, 36, 1, 20, 0)])
def test_lazy_gather_matches_clone_and_index_put(N, R, k, n_tppr, Dm, De, Dt, hot):
    from zebra_b200.embedding import FusedGatherLazy, pending_select
    from tests.test_gpu_gather import _torch_reference
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(N + k)
    E, Dmsg = 800, 2 * Dm + De + Dt
    memory = torch.randn(N, Dm, generator=g).to(dev)
    messages = torch.randn(N, Dmsg, generator=g).to(dev)
    flags = torch.rand(N, generator=g) < 0.6
    ef = torch.randn(E, De, generator=g).to(dev)
    tw = torch.from_numpy(1 / 10 ** np.linspace(0, 9, Dt, dtype=np.float32)).to(dev)
    tb = (torch.randn(Dt, generator=g) * 0.1).to(dev)
    ids = torch.randint(0, hot or N, (n_tppr, R, k), generator=g, dtype=torch.int32)
    if hot:
        ids[:, ::3] = torch.randint(0, N, (n_tppr, (R + 2) // 3, k), generator=g, dtype=torch.int32)
    eidx = torch.randint(0, E, (n_tppr, R, k), generator=g, dtype=torch.int32)
    dt = (torch.rand(n_tppr, R, k, generator=g) * 2.6e6).float()
    w = torch.rand(n_tppr, R, k, generator=g).float()
    w[:, ::7] = 0
    ids[:, ::7] = 0
    w[:, 1::5, k // 2:] = 0
    src = torch.randint(0, N, (R,), generator=g, dtype=torch.int32).to(dev)
    topk = _topk(ids, eidx, dt, w, dev)
    torch.manual_seed(3)
    gru = torch.nn.GRUCell(Dmsg, Dm).to(dev)

    # ours: pending set on the device, updater on its rows only, gather with redirection
    idx, bitmap, wordprefix = pending_select(topk, n_tppr, R, k, flags.to(dev).view(torch.uint8), N)
    upd = gru(messages[idx], memory[idx])
    upd.retain_grad()
    X, wn, srows = FusedGatherLazy.apply(memory, upd, bitmap, wordprefix, topk, src, ef, tw, tb, n_tppr, R, k)
    gX = torch.randn(X.shape, generator=g).to(dev) * (wn.reshape(n_tppr, R * k, 1) != 0)
    gS = torch.randn(srows.shape, generator=g).to(dev)
    (X * gX).sum().add((srows * gS).sum()).backward()
    g_upd = upd.grad.clone()
    g_par = [p.grad.clone() for p in gru.parameters()]
    gru.zero_grad()

    # the reference's sequence (memory_updater.py:61-90 + embedding_module.py:243-264), plain torch
    index = np.unique(ids.numpy())
    uid = index[np.where(flags.numpy()[index])[0]]
    assert np.array_equal(idx.cpu().numpy(), uid)
    updated = memory.clone()
    val = gru(messages[uid], updated[uid])
    val.retain_grad()
    updated[uid] = val
    ref = (updated[src.long()] * gS).sum()
    for j in range(n_tppr):
        Xr, wr, sr = _torch_reference(updated, ef, tw, tb, ids[j].to(dev), eidx[j].to(dev), dt[j].to(dev), w[j].to(dev), src)
        assert torch.equal(X[j], Xr.reshape(R * k, -1)), f"instance {j}: gathered rows differ"
        assert torch.equal(srows, sr)
        ref = ref + (Xr.reshape(R * k, -1) * gX[j]).sum()
    assert torch.equal(upd, val)
    ref.backward()
    assert torch.equal(g_upd, val.grad), f"{int((g_upd != val.grad).sum())} elements of the updater's gradient differ"
    for a, p in zip(g_par, gru.parameters()):
        torch.testing.assert_close(a, p.grad, rtol=1e-5, atol=1e-6)
    from zebra_b200 import _lib
    assert _lib.lib().zebra_gather_check(None) == 0


class _Updater:
    """Stand-in for modules/memory_updater.py: the GRU cell plus get_updated_memory restated from :61-90."""

    def __init__(self, cell):
        self.memory_updater = cell

    def get_updated_memory(self, memory, index=None):
        flags = memory.nodes
        ids = np.where(flags)[0] if index is None else index[np.where(flags[index])[0]]
        if len(ids) == 0:
            return memory.memory.clone(), memory.last_update.clone()
        updated = memory.memory.clone()
        updated[ids] = self.memory_updater(memory.messages[ids], updated[ids])
        last = memory.last_update.clone()
        last[ids] = memory.timestamps[ids]
        return updated, last


def test_level2_embedding_equals_level1_in_training():
    """compute_embedding_tppr_ensemble(train=True) with the lazy update on the device returns the same
    embeddings and hands every parameter the same gradient as the level-1 path, batch after batch."""
    from zebra_b200 import embedding as E
    dev = torch.device("cuda")
    st = synth.make_stream("wikipedia", n_edges=2000, seed=12)
    neg = synth.make_negatives(st, 200)
    N, d_e, Dm = st.shape.table_nodes, 8, 100
    g = torch.Generator(device="cpu").manual_seed(1)
    ef = torch.randn(st.n_edges + 1, d_e, generator=g).to(dev)
    args = types.SimpleNamespace(alpha_list=[0.1, 0.1], beta_list=[0.5, 0.95], topk=20, tppr_strategy="streaming",
                                 n_degree=10, n_layer=2)
    Dmsg = 2 * Dm + d_e + 100
    torch.manual_seed(0)
    mods = [E.GraphDiffusionEmbedding(ef, None, d_e, 100, 100, dev, args, N).to(dev) for _ in range(2)]
    mods[1].load_state_dict(mods[0].state_dict())
    cells = [torch.nn.GRUCell(Dmsg, Dm).to(dev) for _ in range(2)]
    cells[1].load_state_dict(cells[0].state_dict())
    for m in mods:
        m.train()
        for layer in m.modules():
            if isinstance(layer, torch.nn.Dropout):
                layer.p = 0.0                                      # same arithmetic in both arms
    mem = types.SimpleNamespace(memory=torch.randn(N, Dm, generator=g).to(dev), messages=torch.randn(N, Dmsg, generator=g).to(dev),
                                last_update=torch.zeros(N, device=dev), timestamps=torch.zeros(N, device=dev),
                                nodes=np.zeros(N, dtype=bool))
    rng = np.random.default_rng(4)
    old = E.LAZY_MEMORY
    try:
        for b, lo in enumerate(range(0, st.n_edges, 200)):
            hi = lo + 200
            nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]])
            ts = np.concatenate([st.ts[lo:hi]] * 3)
            mem.nodes[:] = rng.random(N) < (0.0 if b == 0 else 0.5)   # first batch: empty pending set
            out = []
            for arm in (0, 1):
                E.LAZY_MEMORY = bool(arm)
                emb = mods[arm].compute_embedding_tppr_ensemble(mem, nodes, ts, st.eidx[lo:hi].astype(np.int64), _Updater(cells[arm]), True)
                for p in list(mods[arm].parameters()) + list(cells[arm].parameters()):
                    p.grad = None
                emb.square().sum().backward()
                out.append((emb.detach(), [None if p.grad is None else p.grad.clone()
                                           for p in list(mods[arm].parameters()) + list(cells[arm].parameters())]))
            assert torch.equal(out[0][0], out[1][0]), f"batch {b}: embeddings differ"
            for x, y in zip(out[0][1], out[1][1]):
                assert (x is None) == (y is None)
                if x is not None:
                    torch.testing.assert_close(x, y, rtol=1e-5, atol=1e-6)
    finally:
        E.LAZY_MEMORY = old


@pytest.mark.parametrize("lazy", [False, True])
def test_train_branch_matches_reference_recording(lazy):
    """Levels 1 and 2 against a recording of the reference's own train branch (its Memory, its
    GRUMemoryUpdater.get_updated_memory, its GraphDiffusionEmbedding; tests/golden/make_golden_embedding_train.py).
    Tolerance 2e-5 / 1e-4 (the recording's GEMMs ran on the CPU)."""
    from tests import helpers as H
    from zebra_b200 import embedding as E
    g = H.load("embedding_train_streaming_wikipedia")
    n_edges, d_e = int(g["n_edges"]), int(g["d_e"])
    st, neg, N, ef, memory = H.embedding_inputs("wikipedia", n_edges, d_e)
    assert N == int(g["num_nodes"]) and np.isclose(float(memory.sum()), float(g["mem_sum"]))
    dev = torch.device("cuda")
    args = types.SimpleNamespace(alpha_list=[0.1, 0.1], beta_list=[0.5, 0.95], topk=20, tppr_strategy="streaming", n_degree=10, n_layer=2)
    mod = E.GraphDiffusionEmbedding(torch.from_numpy(ef).to(dev), None, d_e, 100, 100, dev, args, N).to(dev)
    mod.load_state_dict({k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd.")})
    mod.train()
    mod.drop.p = 0.0
    d_msg = 2 * 100 + d_e + 100
    cell = torch.nn.GRUCell(d_msg, 100).to(dev)
    cell.load_state_dict({k[7:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("gru_sd.")})
    rng = np.random.default_rng(int(g["msg_seed"]))
    messages = (rng.standard_normal((N, d_msg)) * 0.3).astype(np.float32)
    flags = np.unpackbits(g["flags"], axis=1)[:, :N].astype(bool)
    mem = types.SimpleNamespace(memory=torch.from_numpy(memory).to(dev), messages=torch.from_numpy(messages).to(dev),
                                last_update=torch.zeros(N, device=dev), timestamps=torch.zeros(N, device=dev), nodes=None)
    params = {**{f"mod.{k}": v for k, v in mod.named_parameters()}, **{f"gru.memory_updater.{k}": v for k, v in cell.named_parameters()}}
    old = E.LAZY_MEMORY
    E.LAZY_MEMORY = lazy
    try:
        for b, lo in enumerate(range(0, n_edges, 200)):
            hi = lo + 200
            mem.nodes = flags[b].copy()
            nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]])
            ts = np.concatenate([st.ts[lo:hi]] * 3)
            for p in params.values():
                p.grad = None
            emb = mod.compute_embedding_tppr_ensemble(mem, nodes, ts, st.eidx[lo:hi].astype(np.int64), _Updater(cell), True)
            emb.square().sum().backward()
            if f"emb_{b}" in g.files:
                torch.testing.assert_close(emb.detach().cpu(), torch.from_numpy(g[f"emb_{b}"]), rtol=2e-5, atol=2e-5)
                for k in g.files:
                    if k.startswith(f"grad_{b}."):
                        torch.testing.assert_close(params[k[len(f"grad_{b}."):]].grad.cpu(), torch.from_numpy(g[k]), rtol=1e-4, atol=1e-4)
    finally:
        E.LAZY_MEMORY = old
