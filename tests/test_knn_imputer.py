"""KNNImputer(n_neighbors=1) (train_ensemble_public.py:37-40): the CPU restatement against scikit-learn itself, and the
CUDA kernel against the restatement."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import knn_oracle as ko  # noqa: E402


def _cases():
    rng = np.random.default_rng(7)
    A = rng.normal(size=(713, 24))
    A[rng.random(A.shape) < 0.05] = np.nan
    A[:, 7] = np.nan                                   # all-missing column: dropped from the output
    B = rng.normal(size=(300, 24))
    B[rng.random(B.shape) < 0.08] = np.nan
    B[5, :] = np.nan
    B[5, 3] = 0.3                                      # shares one coordinate with the donors
    B[6, :] = np.nan                                   # shares nothing: every column gets the column mean
    yield "continuous", A, B
    A2 = rng.integers(0, 2, size=(500, 20)).astype(float)
    A2[rng.random(A2.shape) < 0.05] = np.nan
    B2 = rng.integers(0, 2, size=(400, 20)).astype(float)
    B2[rng.random(B2.shape) < 0.1] = np.nan
    yield "binary (ties)", A2, B2
    # the candidate matrices the training replay feeds the imputer (oracle/retrain.py; 64 columns, integer-valued: ties)
    from oracle import retrain
    risk = lambda X17: np.full(len(X17), 0.2)         # noqa: E731  (outcomes are irrelevant here)
    dev, _, _ = retrain.synthetic_cohort(2020, 0, 713, risk, missing_rate=0.02)
    sel, _, _ = retrain.synthetic_cohort(2020, 713, 306, risk, missing_rate=0.02)
    yield "replayed candidates (ties)", dev, sel


def test_oracle_matches_scikit_learn_where_it_is_defined():
    from sklearn.impute import KNNImputer
    for name, A, B in _cases():
        imp = KNNImputer(missing_values=np.nan, n_neighbors=1, copy=True).fit(A)
        state = ko.fit(A)
        for Q in (B, A):                               # transform and fit_transform (train:38, train:40)
            want = imp.transform(Q)
            got, cands = ko.transform(state, Q, return_candidates=True)
            assert got.shape == want.shape and not np.isnan(got).any()
            valid_cols = np.flatnonzero(state["valid"])
            pos = {c: k for k, c in enumerate(valid_cols)}
            differing = 0
            for (i, c), values in cands.items():
                # scikit-learn's pick among equidistant donors is undefined; its value must come from one of them
                assert want[i, pos[c]] in values, (name, i, c)
                differing += want[i, pos[c]] != got[i, pos[c]]
            untouched = ~np.isnan(Q[:, valid_cols])
            assert np.array_equal(got[untouched], want[untouched])
            if "ties" not in name:
                assert np.array_equal(got, want), name  # no ties: bit-identical


@pytest.mark.gpu
def test_cuda_imputer_matches_the_oracle_bit_for_bit():
    import mlr_b200 as pkg
    for name, A, B in _cases():
        state = ko.fit(A)
        imp = pkg.KNNImputerB200().fit(A)
        for Q in (B, A, B[:1], B[:0]):
            got = imp.transform(Q)
            want = ko.transform(state, Q)
            assert got.shape == want.shape
            assert np.array_equal(got, want), name     # same tie rule (lowest fit row), same float64 operation order
    with pytest.raises(ValueError):
        pkg.KNNImputerB200(n_neighbors=5)
    with pytest.raises(ValueError):
        pkg.KNNImputerB200().fit(np.zeros((4, 3))).transform(np.zeros((2, 4)))
