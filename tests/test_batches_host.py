"""PassSchedule against a recording of the reference's own evaluation loop head:
slices and seeded negatives of every batch (tests/golden/pass_schedule.npz, made by make_golden_pass.py with the
reference's RandEdgeSampler)."""
import types

import numpy as np
import pytest

from tests import helpers as H
from zebra_b200.batches import PassSchedule, RandEdgeSampler, read_edge_csv


def _split(g):
    lo, hi = int(g["lo"]), int(g["hi"])
    return types.SimpleNamespace(sources=g["src"][lo:hi], destinations=g["dst"][lo:hi], timestamps=g["ts"][lo:hi],
                                 edge_idxs=g["eidx"][lo:hi], n_interactions=hi - lo)


@pytest.mark.parametrize("seed", [0, 3])
@pytest.mark.parametrize("bs", [200, 37])
def test_schedule_draws_the_reference_negatives(seed, bs):
    g = H.load("pass_schedule")
    data = _split(g)
    sampler = RandEdgeSampler(g["src"], g["dst"], seed=seed)
    for _ in range(2):                                  # the reference resets the sampler at the top of every pass
        sampler.reset_random_state()
        ps = PassSchedule(data, sampler, bs)
        assert np.array_equal(ps.negatives, g[f"neg_seed{seed}_bs{bs}"])
    n = data.n_interactions
    assert ps.n_batches == -(-n // bs)
    got = [ps.batch(b) for b in range(ps.n_batches)]
    for q, full in enumerate((data.sources, data.destinations, ps.negatives, data.timestamps, data.edge_idxs)):
        assert np.array_equal(np.concatenate([x[q] for x in got]), full)
    assert all(len(x[0]) == bs for x in got[:-1]) and len(got[-1][0]) == n - bs * (ps.n_batches - 1)


def test_schedule_of_an_empty_split_and_announce_without_a_finder():
    data = types.SimpleNamespace(sources=np.zeros(0, np.int64), destinations=np.zeros(0, np.int64),
                                 timestamps=np.zeros(0), edge_idxs=np.zeros(0, np.int64))
    ps = PassSchedule(data, RandEdgeSampler([1, 2], [3, 4], seed=0), 200)
    assert ps.n_batches == 0 and len(ps.negatives) == 0 and ps.stats() is None
    ps.close()


def test_read_edge_csv(tmp_path):
    p = tmp_path / "ml_x.csv"
    p.write_text(",u,i,ts,label,idx\n0,1,5,0.0,0,1\n1,2,6,3.5,1,2\n2,1,6,9.25,0,3\n")
    u, i, ts, idx, lab = read_edge_csv(str(p))
    assert u.tolist() == [1, 2, 1] and i.tolist() == [5, 6, 6] and ts.tolist() == [0.0, 3.5, 9.25]
    assert idx.tolist() == [1, 2, 3] and lab.tolist() == [0, 1, 0]


def _train(g):
    lo = int(g["lo"])
    return types.SimpleNamespace(sources=g["src"][:lo], destinations=g["dst"][:lo], timestamps=g["ts"][:lo],
                                 edge_idxs=g["eidx"][:lo])


def test_train_lookahead_hands_out_the_reference_draws_and_leaves_the_stream_where_it_was():
    from zebra_b200.batches import TrainLookahead
    g = H.load("pass_schedule")
    data = _train(g)
    n = len(data.sources)
    sampler = RandEdgeSampler(data.sources, data.destinations)
    look = TrainLookahead(data)
    np.random.seed(0)
    for ep in range(2):
        negs = []
        for b in range(-(-n // 200)):
            size = min(n, (b + 1) * 200) - b * 200
            _, neg = look.take(sampler, size, sampler.sample)
            negs.append(neg)
        assert np.array_equal(np.concatenate(negs), g[f"train_neg_epoch{ep}"]), ep
    assert np.array_equal(np.random.randint(0, 1 << 30, 4), g["train_after"])
    assert look.epochs == 2 and look.enabled


def test_train_lookahead_gives_way_to_a_loop_it_did_not_predict():
    from zebra_b200.batches import TrainLookahead
    g = H.load("pass_schedule")
    data = _train(g)
    sizes = [200, 200, 200, 77, 200, 13, 200]
    plain = RandEdgeSampler(data.sources, data.destinations)
    np.random.seed(5)
    want = [plain.sample(s)[1] for s in sizes]
    tail = np.random.randint(0, 1 << 30, 4)
    sampler = RandEdgeSampler(data.sources, data.destinations)
    look = TrainLookahead(data)
    np.random.seed(5)
    got = [look.take(sampler, s, sampler.sample)[1] for s in sizes]
    assert all(np.array_equal(a, b) for a, b in zip(got, want))
    assert np.array_equal(np.random.randint(0, 1 << 30, 4), tail)
    assert not look.enabled
