"""Host restatements of the three scikit-learn metrics the reference calls per batch (zebra_b200/metrics.py;
reference: evaluation/evaluation.py:34-45, train.py:218-227) against scikit-learn itself."""
import numpy as np
import sklearn.metrics as skm

from zebra_b200 import metrics as M


def _batches(rng, count):
    for c in range(count):
        n = int(rng.integers(1, 260))
        if c % 4 == 0:
            pos, neg = rng.random(n), rng.random(n)
        elif c % 4 == 1:                                   # many exactly tied scores
            pos, neg = rng.integers(0, 6, n) / 5.0, rng.integers(0, 6, n) / 5.0
        elif c % 4 == 2:                                   # saturated sigmoid outputs
            pos, neg = 1 / (1 + np.exp(-rng.normal(3, 4, n))), 1 / (1 + np.exp(-rng.normal(-3, 4, n)))
        else:
            pos, neg = np.full(n, 0.5), np.full(n, 0.5)
        yield pos.astype(np.float32).reshape(-1, 1), neg.astype(np.float32).reshape(-1, 1)


def test_host_metrics_equal_sklearn():
    rng = np.random.default_rng(0)
    for pos, neg in _batches(rng, 400):
        size = len(pos)
        score = np.concatenate([pos, neg])                 # [2n,1] exactly as train.py:221 builds it
        label = np.concatenate([np.ones(size), np.zeros(size)])
        zero = np.zeros(size)
        pred = np.argmax(np.hstack([pos, neg]), axis=1)
        assert M.average_precision_score(label, score) == skm.average_precision_score(label, score)
        assert M.roc_auc_score(label, score) == skm.roc_auc_score(label, score)
        assert M.accuracy_score(zero, pred) == skm.accuracy_score(zero, pred)


def test_other_signatures_fall_through_to_sklearn():
    y = np.array([0, 1, 2, 1]); s = np.array([[.1, .7, .2], [.2, .5, .3], [.1, .1, .8], [.3, .4, .3]])
    assert M.roc_auc_score(y, s, multi_class="ovr") == skm.roc_auc_score(y, s, multi_class="ovr")
    assert M.accuracy_score(np.array([1, 0, 1]), np.array([1, 1, 1]), normalize=False) == 2
