"""zebra_edge_metrics (csrc/metrics_api.cu) against scikit-learn on the batches the reference scores
(evaluation/evaluation.py:34-45): average precision, ROC AUC, accuracy.  Tolerance 1e-12 absolute (the kernel
sums the same terms in another order than scikit-learn's cumsum / diff formulation)."""
import numpy as np
import pytest
import sklearn.metrics as skm
import torch

from tests.test_metrics_host import _batches

pytestmark = pytest.mark.gpu


def test_edge_metrics_kernel_matches_sklearn():
    from zebra_b200.metrics import edge_metrics
    rng = np.random.default_rng(1)
    for pos, neg in _batches(rng, 120):
        got = edge_metrics(torch.from_numpy(pos).cuda(), torch.from_numpy(neg).cuda()).cpu().numpy()[0]
        size = len(pos)
        score = np.concatenate([pos, neg])
        label = np.concatenate([np.ones(size), np.zeros(size)])
        if size >= 1:
            assert abs(got[0] - skm.average_precision_score(label, score)) < 1e-12
            assert abs(got[1] - skm.roc_auc_score(label, score)) < 1e-12
            assert abs(got[2] - skm.accuracy_score(np.zeros(size), np.argmax(np.hstack([pos, neg]), axis=1))) < 1e-12


def test_edge_metrics_many_batches_in_one_launch():
    from zebra_b200.metrics import edge_metrics
    rng = np.random.default_rng(2)
    n, bs = 200 * 37 + 61, 200                        # ragged last batch
    pos = rng.random(n).astype(np.float32); neg = (rng.random(n) * 0.9).astype(np.float32)
    got = edge_metrics(torch.from_numpy(pos).cuda(), torch.from_numpy(neg).cuda(), batch_size=bs).cpu().numpy()
    assert got.shape == (38, 3)
    for b in range(38):
        p, q = pos[b * bs:(b + 1) * bs], neg[b * bs:(b + 1) * bs]
        label = np.concatenate([np.ones(len(p)), np.zeros(len(q))])
        assert abs(got[b, 0] - skm.average_precision_score(label, np.concatenate([p, q]))) < 1e-12
        assert abs(got[b, 1] - skm.roc_auc_score(label, np.concatenate([p, q]))) < 1e-12
        assert abs(got[b, 2] - np.mean(p >= q)) < 1e-12
