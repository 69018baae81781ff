"""Arrays -> live scikit-learn 1.9 estimators, so scikit-learn's own C++/Cython scores a given ``ModelArrays``.

TEST INFRASTRUCTURE ONLY (see oracle/stack_oracle.py header).  This is how the *published* parameters (statically
decoded, never unpickled) and synthetic tree tables get scored by the reference's real arithmetic: every fitted
attribute is assigned by hand on freshly constructed estimators, trees go through ``Tree.__setstate__``
(SK/tree/_tree.pyx:819-848) with node records in the 1.9 8-field layout (SK/tree/_tree.pxd:15-25).
"""
from __future__ import annotations

import warnings

import numpy as np


def _tree_from_table(t, n_features):
    from sklearn.tree import DecisionTreeRegressor
    from sklearn.tree._tree import Tree, NODE_DTYPE  # noqa: N811

    n = t.left.shape[0]
    nodes = np.zeros(n, dtype=NODE_DTYPE)
    nodes["left_child"] = t.left
    nodes["right_child"] = t.right
    nodes["feature"] = t.feature
    nodes["threshold"] = t.threshold
    nodes["impurity"] = 0.0
    nodes["n_node_samples"] = 1
    nodes["weighted_n_node_samples"] = 1.0
    if "missing_go_to_left" in NODE_DTYPE.names:
        nodes["missing_go_to_left"] = 0
    depth = np.zeros(n, dtype=np.int64)
    for i in range(n):
        if t.left[i] != -1:
            depth[t.left[i]] = depth[i] + 1
            depth[t.right[i]] = depth[i] + 1
    tree = Tree(n_features, np.array([1], dtype=np.intp), 1)
    tree.__setstate__({"max_depth": int(depth.max()), "node_count": n, "nodes": nodes,
                       "values": np.ascontiguousarray(t.value.reshape(n, 1, 1), dtype=np.float64)})
    reg = DecisionTreeRegressor()
    reg.tree_ = tree
    reg.n_features_in_ = n_features
    reg.n_outputs_ = 1
    reg.max_features_ = n_features
    return reg


def _rebuild_svc(m, n_features):
    from sklearn.pipeline import Pipeline
    from sklearn.preprocessing import StandardScaler
    from sklearn.svm import SVC

    sc = StandardScaler()
    sc.mean_ = m.mean.copy()
    sc.scale_ = m.scale.copy()
    sc.var_ = m.scale ** 2
    sc.n_features_in_ = n_features
    sc.n_samples_seen_ = 1
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        svc = SVC(kernel="rbf", probability=True, gamma=m.gamma)
    n_sv = m.sv.shape[0]
    svc._sparse = False
    svc.support_ = np.arange(n_sv, dtype=np.int32)
    svc.support_vectors_ = np.ascontiguousarray(m.sv, dtype=np.float64)
    svc._n_support = np.asarray(m.n_support, dtype=np.int32)
    svc._dual_coef_ = np.ascontiguousarray(m.dual_coef.reshape(1, n_sv))
    svc.dual_coef_ = -svc._dual_coef_
    svc._intercept_ = np.array([m.intercept])
    svc.intercept_ = -svc._intercept_
    svc._probA = np.array([m.prob_a])
    svc._probB = np.array([m.prob_b])
    svc._gamma = float(m.gamma)
    svc.classes_ = np.array([0, 1])
    svc.class_weight_ = np.ones(2)
    svc.shape_fit_ = (n_sv, n_features)
    svc.fit_status_ = 0
    svc.n_features_in_ = n_features
    svc.n_iter_ = np.array([1], dtype=np.int32)
    svc._effective_probability = True
    return Pipeline([("standardscaler", sc), ("svc", svc)])


def _rebuild_gbt(m, n_features):
    from sklearn.dummy import DummyClassifier
    from sklearn.ensemble import GradientBoostingClassifier
    from sklearn._loss.loss import HalfBinomialLoss

    g = GradientBoostingClassifier(n_estimators=len(m.trees), learning_rate=m.learning_rate)
    g.n_estimators_ = len(m.trees)
    g.n_features_in_ = n_features
    g.n_classes_ = 2
    g.n_trees_per_iteration_ = 1
    g.classes_ = np.array([0, 1])
    g.max_features_ = n_features
    g._loss = HalfBinomialLoss(sample_weight=None)
    p1 = 1.0 / (1.0 + np.exp(-m.init_raw))
    init = DummyClassifier(strategy="prior")
    init.classes_ = np.array([0, 1])
    init.n_classes_ = 2
    init.class_prior_ = np.array([1.0 - p1, p1])
    init.n_outputs_ = 1
    init.sparse_output_ = False
    init._strategy = "prior"
    init.n_features_in_ = n_features
    g.init_ = init
    ests = np.empty((len(m.trees), 1), dtype=object)
    for i, t in enumerate(m.trees):
        ests[i, 0] = _tree_from_table(t, n_features)
    g.estimators_ = ests
    g.train_score_ = np.zeros(len(m.trees))
    return g


def _rebuild_lr(coef, intercept):
    from sklearn.linear_model import LogisticRegression

    lr = LogisticRegression()
    lr.coef_ = np.ascontiguousarray(coef.reshape(1, -1), dtype=np.float64)
    lr.intercept_ = np.array([intercept], dtype=np.float64)
    lr.classes_ = np.array([0, 1])      # members are fitted on label-encoded y (SK/ensemble/_stacking.py fit)
    lr.n_features_in_ = coef.shape[0]
    lr.n_iter_ = np.array([1], dtype=np.int32)
    return lr


def rebuild_stack(model):
    """``ModelArrays`` -> fitted-looking ``StackingClassifier`` whose predict_proba runs sklearn's own code."""
    from sklearn.ensemble import StackingClassifier
    from sklearn.preprocessing import LabelEncoder

    members = []
    for m in model.members:
        if m.kind == "svc":
            est = _rebuild_svc(m, model.n_features)
        elif m.kind == "gbt":
            est = _rebuild_gbt(m, model.n_features)
        else:
            est = _rebuild_lr(m.coef, m.intercept)
        members.append((m.name, est))
    final = _rebuild_lr(model.final_coef, model.final_intercept)
    st = StackingClassifier(estimators=members, final_estimator=final)
    st.estimators_ = [e for _, e in members]
    st.named_estimators_ = dict(members)
    st.final_estimator_ = final
    st.stack_method_ = ["predict_proba"] * len(members)
    st.classes_ = np.asarray(model.classes, dtype=np.float64)
    le = LabelEncoder()
    le.classes_ = st.classes_
    st._le = le  # older attribute name, harmless
    st._label_encoder = le
    return st
