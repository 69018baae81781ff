"""Headless replay of the reference's model-development recipe on a synthetic cohort.  TEST INFRASTRUCTURE ONLY.

train_ensemble_public.py cannot be imported in this image (it imports matplotlib at line 2 and calls plotting helpers
that scikit-learn removed in 1.2 at lines 67 and 79), and its two ``.mat`` inputs are not published.  This module
writes synthetic ``develop_data.mat`` / ``model_select_data.mat`` files that satisfy load_data_public.py:4-14 (keys
``data_tb`` with the outcome in the last column and ``clin_var_names``), loads them back through the reference's own
``load_data`` when /root/reference is mounted (else through the product's equivalent loader), and then performs the
steps of train_ensemble_public.py:36-63 with scikit-learn: 1-NN imputation, LassoCV -> SelectFromModel(max 17
features), the three-member stack and its fit.  Lines 66-90 (plots) are left out.
"""
from __future__ import annotations

import importlib
import os
import sys
import warnings

import numpy as np

_pkg = importlib.import_module("machine-learning-replications_b200")
synth = importlib.import_module("machine-learning-replications_b200.synth")

REFERENCE_DIR = "/root/reference/Machine Learning for Predicting Heart Failure Progression"
N_DEVELOP = 713      # StandardScaler.n_samples_seen_ of the published model
N_SELECT = 714       # Table 1's 1427 patients minus the development rows
INIT_RS = 2020       # train_ensemble_public.py:30


def _uniform(seed, rows, stream):
    return synth.hash_u32(seed, rows, 200 + stream).astype(np.float64) / 2.0 ** 32


def synthetic_cohort(seed, row0, n, risk_fn, missing_rate=0.004):
    """[n, 64] float64 candidate matrix (a few NaNs, as the imputer at train:37-40 expects) and outcomes [n]."""
    specs = synth.candidate_specs()
    rows = np.arange(row0, row0 + n, dtype=np.uint64)
    X = synth.host_rows(specs, seed, row0, n).astype(np.float64)
    y = (_uniform(seed, rows, 1) < risk_fn(X[:, :17])).astype(np.float64)
    # knock out a few non-schema entries so KNNImputer has something to fill
    for j in range(17, synth.N_CANDIDATES, 5):
        miss = synth.hash_u32(seed + 1, rows, j) < int(missing_rate * 2 ** 32)
        X[miss, j] = np.nan
    names = np.array([[s.name for s in specs]], dtype=object)
    return X, y, names


def write_mat(path, X, y, names):
    import scipy.io as sio
    sio.savemat(path, {"data_tb": np.column_stack([X, y]), "clin_var_names": names})


def load_cohort(path):
    """The reference's loader when it is mounted, the product's equivalent otherwise."""
    if os.path.isdir(REFERENCE_DIR):
        sys.path.insert(0, REFERENCE_DIR)
        try:
            from load_data_public import load_data  # type: ignore
            return load_data(path)
        finally:
            sys.path.remove(REFERENCE_DIR)
    return _pkg.load_cohort_mat(path)


def replay_training(workdir, risk_fn, seed=INIT_RS, verbose=False):
    """Returns ``(clf, support_mask, X_sel_optm, y_sel, feature_names)`` following train:36-63."""
    from sklearn.ensemble import GradientBoostingClassifier, StackingClassifier
    from sklearn.feature_selection import SelectFromModel
    from sklearn.impute import KNNImputer
    from sklearn.linear_model import LassoCV, LogisticRegression
    from sklearn.pipeline import make_pipeline
    from sklearn.preprocessing import StandardScaler
    from sklearn.svm import SVC

    os.makedirs(workdir, exist_ok=True)
    dev_path = os.path.join(workdir, "develop_data.mat")
    sel_path = os.path.join(workdir, "model_select_data.mat")
    write_mat(dev_path, *synthetic_cohort(seed, 0, N_DEVELOP, risk_fn))
    write_mat(sel_path, *synthetic_cohort(seed, N_DEVELOP, N_SELECT, risk_fn))

    np.random.seed(seed)                                            # train:31
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")                             # FutureWarnings for penalty= / probability=
        X_dev, y_dev, names = load_cohort(dev_path)                 # train:36
        imputer = KNNImputer(missing_values=np.nan, n_neighbors=1, copy=True)
        X_dev = imputer.fit_transform(X_dev)                        # train:37-38
        X_sel, y_sel, _ = load_cohort(sel_path)                     # train:39
        X_sel = imputer.transform(X_sel)                            # train:40
        members = [                                                 # train:43-47
            ("svc", make_pipeline(StandardScaler(), SVC(class_weight="balanced", probability=True,
                                                        random_state=seed))),
            ("gbc", GradientBoostingClassifier(n_estimators=100, max_depth=1, random_state=seed)),
            ("lg", LogisticRegression(class_weight="balanced", penalty="l1", solver="liblinear")),
        ]
        clf = StackingClassifier(estimators=members,
                                 final_estimator=LogisticRegression(class_weight="balanced"))   # train:48
        lasso = LassoCV(random_state=seed, cv=10)                   # train:51
        sfm = SelectFromModel(lasso, threshold=-np.inf, max_features=17).fit(X_dev, y_dev)       # train:52
        mask = sfm.get_support()
        clf.fit(X_dev[:, mask], y_dev)                              # train:53,61
    if verbose:
        print("selected columns:", names[0, mask])
    return clf, mask, X_sel[:, mask], y_sel, names
