"""``KNNImputer(missing_values=np.nan, n_neighbors=1)`` on the GPU -- the preprocessing step of
train_ensemble_public.py:37-40, applied to the candidate matrix before feature selection.

Same estimator protocol (``fit`` / ``transform`` / ``fit_transform``), float64 in and out, all-missing columns dropped
like scikit-learn does (``keep_empty_features=False``).  Distances and donor values are float64, in scikit-learn's
operation order (SK/metrics/pairwise.py ``nan_euclidean_distances``, SK/impute/_knn.py).  Among equidistant donors
scikit-learn's pick is an implementation detail of ``np.argpartition``; here the lowest fit row wins.
"""
from __future__ import annotations

import numpy as np

from . import capi


class KNNImputerB200:
    def __init__(self, missing_values=np.nan, n_neighbors=1, copy=True, device=0):
        if not (isinstance(missing_values, float) and np.isnan(missing_values)):
            raise ValueError("only missing_values=np.nan is supported")
        if n_neighbors != 1:
            raise ValueError("only n_neighbors=1 (the reference's setting) is supported")
        self.missing_values, self.n_neighbors, self.copy, self.device = missing_values, 1, copy, int(device)
        self._lib = capi.load_library()

    @staticmethod
    def _check(X, n_features=None):
        X = np.array(X, dtype=np.float64, order="C", copy=True)
        if X.ndim != 2:
            raise ValueError(f"Expected 2D array, got {X.ndim}D array instead")
        if np.isinf(X).any():
            raise ValueError("Input X contains infinity or a value too large for dtype('float64').")
        if n_features is not None and X.shape[1] != n_features:
            raise ValueError(f"X has {X.shape[1]} features, but KNNImputerB200 is expecting {n_features} features as input.")
        return X

    def fit(self, X, y=None):
        import torch
        X = self._check(X)
        mask = np.isnan(X)
        self._valid_mask = ~mask.all(axis=0)
        self.n_features_in_ = X.shape[1]
        # the masked mean scikit-learn takes (SK/impute/_knn.py): sum over the zero-filled column / present count
        mean = np.array([float(np.ma.array(X[:, c], mask=mask[:, c]).mean()) if self._valid_mask[c] else np.nan
                         for c in range(X.shape[1])])
        dev = f"cuda:{self.device}"
        self._fit_X = torch.from_numpy(X).to(dev)
        self._col_mean = torch.from_numpy(mean).to(dev)
        return self

    def transform(self, X):
        import torch
        X = self._check(X, self.n_features_in_)
        n = X.shape[0]
        if n:
            with torch.cuda.device(self.device):
                x = torch.from_numpy(X).to(self._fit_X.device)
                capi.check(self._lib.hfe_knn1_impute(self._fit_X.data_ptr(), self._fit_X.shape[0], self.n_features_in_,
                                                     self._col_mean.data_ptr(), x.data_ptr(), n,
                                                     torch.cuda.current_stream(self.device).cuda_stream))
                X = x.cpu().numpy()
        return X[:, self._valid_mask]

    def fit_transform(self, X, y=None):
        return self.fit(X).transform(X)
