"""Evaluation of the probabilities on the device: what train_ensemble_public.py does right after scoring.

``train_ensemble_public.py:63-64`` thresholds ``y_pred[:,1] > 0.5`` and prints ``metrics.classification_report``;
lines 67 and 79 draw the ROC and precision-recall curves.  ``binary_report`` reproduces the report's numbers from a
confusion matrix counted on the GPU (exact integer counts); ``roc_curve`` / ``roc_auc`` / ``average_precision`` come
from per-class histograms of P(class 1) with 65 536 bins (a binned curve: two probabilities closer than 1.5e-5 share a
threshold).  Plotting itself is out of scope (matplotlib is not installed; the plot helpers left scikit-learn in 1.2).
"""
from __future__ import annotations

import numpy as np

from . import capi


def device_counts(proba, y, threshold=0.5, want_hist=True):
    """``proba``: float32 CUDA tensor ``[n,2]``; ``y``: uint8 CUDA tensor of class indices.  Returns
    ``(confusion [2,2] int64 as [[tn, fp],[fn, tp]], hist [2, BINS] int64 or None)`` on the host."""
    import torch

    lib = capi.load_library()
    n = int(proba.shape[0])
    dev = proba.device
    conf = torch.zeros(4, dtype=torch.int64, device=dev)
    hist = torch.zeros((2, capi.METRIC_BINS), dtype=torch.int32, device=dev) if want_hist else None
    with torch.cuda.device(dev):            # the entry point runs on the current device: make it the tensors' device
        capi.check(lib.hfe_binary_metrics(proba.data_ptr(), y.data_ptr(), n, float(threshold), conf.data_ptr(),
                                          hist.data_ptr() if want_hist else None,
                                          torch.cuda.current_stream(dev).cuda_stream))
    return conf.cpu().numpy().reshape(2, 2), (hist.cpu().numpy().astype(np.int64) if want_hist else None)


def binary_report(confusion):
    """The numbers of ``sklearn.metrics.classification_report(y_true, y_pred, output_dict=True)`` for two classes."""
    (tn, fp), (fn, tp) = np.asarray(confusion, dtype=np.float64)

    def prf(tp_, fp_, fn_):
        p = tp_ / (tp_ + fp_) if tp_ + fp_ else 0.0
        r = tp_ / (tp_ + fn_) if tp_ + fn_ else 0.0
        f = 2 * p * r / (p + r) if p + r else 0.0
        return p, r, f

    p0, r0, f0 = prf(tn, fn, fp)
    p1, r1, f1 = prf(tp, fp, fn)
    s0, s1 = tn + fp, fn + tp
    total = s0 + s1
    return {
        "0": {"precision": p0, "recall": r0, "f1-score": f0, "support": s0},
        "1": {"precision": p1, "recall": r1, "f1-score": f1, "support": s1},
        "accuracy": (tn + tp) / total if total else 0.0,
        "macro avg": {"precision": (p0 + p1) / 2, "recall": (r0 + r1) / 2, "f1-score": (f0 + f1) / 2, "support": total},
        "weighted avg": {"precision": (p0 * s0 + p1 * s1) / total, "recall": (r0 * s0 + r1 * s1) / total,
                         "f1-score": (f0 * s0 + f1 * s1) / total, "support": total},
    }


def roc_curve(hist):
    """(fpr, tpr, thresholds) with thresholds descending, one point per occupied bin (plus the origin)."""
    neg, pos = hist[0][::-1], hist[1][::-1]
    keep = (neg + pos) > 0
    tps = np.cumsum(pos)[keep]
    fps = np.cumsum(neg)[keep]
    thr = (np.arange(hist.shape[1])[::-1] / (hist.shape[1] - 1.0))[keep]
    tpr = np.concatenate([[0.0], tps / max(tps[-1], 1)])
    fpr = np.concatenate([[0.0], fps / max(fps[-1], 1)])
    return fpr, tpr, np.concatenate([[np.inf], thr])


def roc_auc(hist):
    fpr, tpr, _ = roc_curve(hist)
    return float(np.trapezoid(tpr, fpr)) if hasattr(np, "trapezoid") else float(np.trapz(tpr, fpr))


def average_precision(hist):
    """Step-wise area under the precision-recall curve (sklearn.metrics.average_precision_score's definition)."""
    neg, pos = hist[0][::-1], hist[1][::-1]
    keep = (neg + pos) > 0
    tps = np.cumsum(pos)[keep].astype(np.float64)
    fps = np.cumsum(neg)[keep].astype(np.float64)
    precision = tps / np.maximum(tps + fps, 1)
    recall = tps / max(tps[-1], 1)
    return float(np.sum(np.diff(np.concatenate([[0.0], recall])) * precision))
