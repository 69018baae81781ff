"""CPU restatement of ``KNNImputer(missing_values=np.nan, n_neighbors=1)`` as train_ensemble_public.py:37-40 uses it.

Test infrastructure only: nothing under ``machine-learning-replications_b200/`` imports this file.

Follows scikit-learn 1.9.0:
  * ``SK/impute/_knn.py`` ``KNNImputer.fit`` (keeps ``_fit_X``, its missing mask, ``_valid_mask`` = columns that are not
    all-missing) and ``transform`` (per receiver row and missing column: donors = fit rows where that column is present,
    nearest donor by nan-euclidean distance, column mean when every distance is NaN, all-missing columns dropped);
  * ``SK/metrics/pairwise.py`` ``nan_euclidean_distances``: ``sqrt(n_features / present * sum_present (x - y)^2)``, NaN
    when the two rows share no present coordinate.  scikit-learn evaluates the sum through |x|^2 - 2xy + |y|^2 with
    BLAS; here it is the direct sum, which is the same number for integer-valued cohorts and differs by rounding
    (~1e-16 relative) otherwise.

One thing is *not* defined by the reference: which donor wins among equidistant ones.  scikit-learn asks
``np.argpartition(dist, 0)``, whose pick among ties is an implementation detail of numpy's selection kernel (on this
image's AVX-512 build it is not the first minimum).  This restatement and the CUDA kernel take the lowest donor index;
tests compare with scikit-learn exactly where distances have no ties and otherwise check that scikit-learn's value comes
from *a* nearest donor.
"""
import numpy as np


def fit(fit_X):
    fit_X = np.array(fit_X, dtype=np.float64, copy=True)
    mask = np.isnan(fit_X)
    valid = ~mask.all(axis=0)
    # SK/impute/_knn.py: np.ma.array(fit_X[:, col], mask=mask_fit_X[:, col]).mean() -- the masked sum runs over the
    # zero-filled column, so its rounding differs from a mean over the compressed values
    col_mean = np.array([float(np.ma.array(fit_X[:, c], mask=mask[:, c]).mean()) if valid[c] else np.nan
                         for c in range(fit_X.shape[1])])
    return {"fit_X": fit_X, "mask": mask, "valid": valid, "col_mean": col_mean}


def distances(state, x):
    """nan-euclidean distance of one row to every fit row (NaN where nothing is shared)."""
    Y, my = state["fit_X"], state["mask"]
    mx = np.isnan(x)
    present = ~mx[None, :] & ~my
    diff = np.where(present, x[None, :] - Y, 0.0)
    s = (diff * diff).sum(axis=1)
    cnt = present.sum(axis=1)
    with np.errstate(invalid="ignore", divide="ignore"):
        d = np.sqrt(s / np.maximum(cnt, 1) * Y.shape[1])
    d[cnt == 0] = np.nan
    return d


def transform(state, X, return_candidates=False):
    """Imputed copy of X without the all-missing columns.  With return_candidates also a dict
    (row, col) -> set of values held by the equidistant nearest donors (for the tie check against scikit-learn)."""
    X = np.array(X, dtype=np.float64, copy=True)
    Y, my, valid = state["fit_X"], state["mask"], state["valid"]
    cands = {}
    for i in range(X.shape[0]):
        miss = np.isnan(X[i]) & valid
        if not miss.any():
            continue
        d = distances(state, X[i])
        for c in np.flatnonzero(miss):
            donors = np.flatnonzero(~my[:, c])
            dd = d[donors]
            if np.isnan(dd).all():
                X[i, c] = state["col_mean"][c]
                continue
            best = np.nanmin(dd)
            ties = donors[dd == best]
            X[i, c] = Y[ties[0], c]                       # lowest donor index among the nearest
            if return_candidates:
                cands[(i, c)] = set(Y[ties, c].tolist())
    out = X[:, valid]
    return (out, cands) if return_candidates else out
