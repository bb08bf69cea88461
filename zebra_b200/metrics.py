"""Link-prediction metrics of the reference's loops without scikit-learn on the critical path (SURVEY.md
section 8 row f4; reference: evaluation/evaluation.py:7-47, train.py:218-227).

* `edge_metrics(pos, neg, batch_size)` — CUDA kernel (csrc/metrics_api.cu): average precision, ROC AUC and
  accuracy of every batch from the probabilities where they are (device tensors), into a device tensor [B, 3].
* `eval_edge_prediction(model, negative_edge_sampler, data, n_neighbors, batch_size)` — same signature and
  return value as the reference's function; the per-batch `.cpu()` synchronisation and the three scikit-learn
  calls are replaced by the kernel and ONE device-to-host copy at the end of the pass.
* `average_precision_score / roc_auc_score / accuracy_score` — host restatements of the scikit-learn
  definitions for the 1-D binary case the reference uses (train.py:218-227 has the numbers on the host
  already); anything else is passed through to scikit-learn.  ~20x cheaper per call than scikit-learn's
  validated entry points on 400 scores.
`install()` (launcher level 3) puts them where the reference's unmodified train.py picks them up.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np

PASS_SCHEDULE = os.environ.get("ZEBRA_PASS_SCHEDULE", "1") != "0"    # announce validation / test passes to the finder
PASS_WINDOW = int(os.environ.get("ZEBRA_PASS_WINDOW", "64"))          # batches streamed per launch, ahead of the model
LAST_PASS = None                                                      # PassSchedule.stats() of the last pass (diagnostics)


def edge_metrics(pos, neg, batch_size=None):
    """pos, neg: float32 CUDA tensors [n] (or [n,1]); returns a float64 CUDA tensor [ceil(n/batch_size), 3] =
    (average precision, ROC AUC, accuracy) per batch.  Asynchronous on the current stream."""
    import torch
    from ._lib import check, lib
    pos = pos.reshape(-1).contiguous().float()
    neg = neg.reshape(-1).contiguous().float()
    n = pos.numel()
    if neg.numel() != n:
        raise ValueError("pos and neg must have the same length")
    bs = int(batch_size or n)
    nb = (n + bs - 1) // bs if n else 0
    out = torch.empty((nb, 3), dtype=torch.float64, device=pos.device)
    if n:
        with torch.cuda.device(pos.device):
            check(lib().zebra_edge_metrics(C.c_void_p(pos.data_ptr()), C.c_void_p(neg.data_ptr()), n, bs,
                                           C.c_void_p(out.data_ptr()), C.c_void_p(torch.cuda.current_stream(pos.device).cuda_stream)))
    return out


def eval_edge_prediction(model, negative_edge_sampler, data, n_neighbors, batch_size):
    """Drop-in for evaluation/evaluation.py:7-47: identical batches, identical seeded negatives, identical
    model calls; the probabilities stay on the device and are scored there."""
    import torch
    from .batches import PassSchedule
    assert negative_edge_sampler.seed is not None
    negative_edge_sampler.reset_random_state()
    rows = []
    with torch.no_grad():
        model = model.eval()
        # the whole pass assembled before the loop (row f2): same slices, same seeded draws in the same order;
        # the streaming finder is told what is coming and computes the T-PPR rows a window of batches ahead
        sched = PassSchedule(data, negative_edge_sampler, batch_size)
        if PASS_SCHEDULE:
            sched.announce(model, PASS_WINDOW)
        try:
            for batch_idx in range(sched.n_batches):
                src, dst, negative_samples, ts, eidx = sched.batch(batch_idx)
                pos_prob, neg_prob = model.compute_edge_probabilities(src, dst, negative_samples, ts, eidx, n_neighbors,
                                                                      train=False)
                rows.append(edge_metrics(pos_prob, neg_prob))
        finally:
            sched.close()
            global LAST_PASS
            LAST_PASS = sched.stats()
    if not rows:
        return float("nan"), float("nan"), float("nan")
    m = torch.cat(rows).cpu().numpy()            # the one device-to-host copy of the pass
    return np.mean(m[:, 0]), np.mean(m[:, 1]), np.mean(m[:, 2])


# ---- host restatements (binary labels, 1-D scores) ----------------------------------------------------

def _binary_curve(y_true, y_score):
    """Cumulative true / false positives at the distinct score values, descending (scikit-learn's
    _binary_clf_curve: stable descending sort, thresholds where the score changes)."""
    order = np.argsort(y_score, kind="mergesort")[::-1]
    ys, yt = y_score[order], y_true[order]
    idx = np.r_[np.where(np.diff(ys))[0], yt.size - 1]
    tps = np.cumsum(yt, dtype=np.float64)[idx]
    fps = 1 + idx - tps
    return fps, tps


def _simple(y_true, y_score, kw):
    if kw:
        return None
    yt, ys = np.asarray(y_true), np.asarray(y_score)
    if ys.ndim == 2 and ys.shape[1] == 1:
        ys = ys[:, 0]
    if yt.ndim != 1 or ys.ndim != 1 or yt.shape != ys.shape or yt.size < 2:
        return None
    if not np.all((yt == 0) | (yt == 1)) or yt.min() == yt.max():
        return None
    return yt.astype(np.float64), ys.astype(np.float64)


def average_precision_score(y_true, y_score, **kw):
    a = _simple(y_true, y_score, kw)
    if a is None:
        return _sk("average_precision_score")(y_true, y_score, **kw)
    fps, tps = _binary_curve(*a)
    ps = tps + fps
    precision = np.zeros_like(tps)
    np.divide(tps, ps, out=precision, where=ps != 0)
    recall = tps / tps[-1]
    precision, recall = np.hstack((precision[::-1], 1.0)), np.hstack((recall[::-1], 0.0))
    return float(max(0.0, -np.sum(np.diff(recall) * precision[:-1])))


def roc_auc_score(y_true, y_score, **kw):
    a = _simple(y_true, y_score, kw)
    if a is None:
        return _sk("roc_auc_score")(y_true, y_score, **kw)
    fps, tps = _binary_curve(*a)
    if len(fps) > 2:                                  # scikit-learn drops collinear points (drop_intermediate=True)
        keep = np.where(np.r_[True, np.logical_or(np.diff(fps, 2), np.diff(tps, 2)), True])[0]
        fps, tps = fps[keep], tps[keep]
    tps, fps = np.r_[0.0, tps], np.r_[0.0, fps]
    fpr, tpr = fps / fps[-1], tps / tps[-1]
    return float(np.trapezoid(tpr, fpr)) if hasattr(np, "trapezoid") else float(np.trapz(tpr, fpr))


def accuracy_score(y_true, y_pred, **kw):
    if kw:
        return _sk("accuracy_score")(y_true, y_pred, **kw)
    yt, yp = np.asarray(y_true), np.asarray(y_pred)
    if yt.ndim != 1 or yt.shape != yp.shape:
        return _sk("accuracy_score")(y_true, y_pred)
    return float(np.mean(yt == yp))


_SK = {}


def install():
    """Launcher level 3: scikit-learn's three entry points -> the host restatements (picked up by the
    reference's `from sklearn.metrics import ...`), the reference's eval_edge_prediction -> the device path."""
    import sklearn.metrics as skm
    if not _SK:
        for name in ("average_precision_score", "roc_auc_score", "accuracy_score"):
            _SK[name] = getattr(skm, name)
    skm.average_precision_score = average_precision_score
    skm.roc_auc_score = roc_auc_score
    skm.accuracy_score = accuracy_score
    import evaluation.evaluation as ev                # the reference's module (reference root is on sys.path)
    for name in ("average_precision_score", "roc_auc_score", "accuracy_score"):
        setattr(ev, name, globals()[name])
    ev.eval_edge_prediction = eval_edge_prediction


def _sk(name):
    if not _SK:
        import sklearn.metrics as skm
        for nm in ("average_precision_score", "roc_auc_score", "accuracy_score"):
            _SK[nm] = getattr(skm, nm)
    return _SK[name]
