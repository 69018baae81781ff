"""Attribute walker: fitted stack  ->  plain arrays (``ModelArrays``).

Replaces the reference's ``pickle.load`` of ``hf_predict_model.pkl`` (predict_hf.py:33-34) as the way model
parameters reach the scorer.  Two sources produce the same ``ModelArrays``:

* a *live* scikit-learn ``StackingClassifier`` fitted the way train_ensemble_public.py:43-48,61 does it, and
* the inert tree that ``pickle_static.decode_pickle_file`` makes out of the published pickle.

Only fitted attributes are read; no estimator method is ever called.  The stack must look like the reference's: each
member is one of ``Pipeline(StandardScaler, SVC(kernel='rbf', probability=True))``, a binary
``GradientBoostingClassifier`` (log-loss, K=1) or a binary ``LogisticRegression``; every member is consumed through
``predict_proba`` (``stack_method_``), ``passthrough`` is off, and the combiner is a binary ``LogisticRegression``.
Anything else raises ``ModelShapeError`` instead of being scored approximately.

Sign convention for the SVC member: libsvm is handed the *underscore* attributes ``_dual_coef_`` / ``_intercept_``
(sklearn/svm/_base.py:943-968), the public ones are their negation for binary problems (``_base.py:301-305``); the
decision value libsvm turns into a probability is ``sum_j _dual_coef_[j]*k_j + _intercept_`` because
``model->rho = -intercept`` (libsvm_helper.c:171) and ``predict_values`` subtracts ``rho`` (svm.cpp:2821-2904).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from .pickle_static import Inert, unwrap


class ModelShapeError(ValueError):
    """The fitted object is not a stack this scorer implements."""


@dataclass
class SvcMember:
    """StandardScaler -> RBF SVC with libsvm probability outputs."""
    name: str
    mean: np.ndarray          # [F] f64 scaler mean_
    scale: np.ndarray         # [F] f64 scaler scale_
    gamma: float              # _gamma
    sv: np.ndarray            # [n_sv, F] f64, in scaled space
    dual_coef: np.ndarray     # [n_sv] f64 == _dual_coef_[0]
    intercept: float          # _intercept_[0]  (== -rho)
    prob_a: float             # _probA[0]
    prob_b: float             # _probB[0]
    n_support: np.ndarray     # [2] int64
    kind: str = "svc"


@dataclass
class TreeTable:
    """One regression tree of the boosting member, as sklearn stores it (tree/_tree.pxd:15-25)."""
    left: np.ndarray          # [n_nodes] int32, -1 at leaves
    right: np.ndarray         # [n_nodes] int32
    feature: np.ndarray       # [n_nodes] int32, -2 at leaves
    threshold: np.ndarray     # [n_nodes] f64
    value: np.ndarray         # [n_nodes] f64 (values[:,0,0])


@dataclass
class GbtMember:
    """Binary gradient boosting, K=1 (rows a8/a9): raw = init_raw + learning_rate * sum_t value_t[leaf_t(x)]."""
    name: str
    init_raw: float
    learning_rate: float
    trees: list               # list[TreeTable], boosting order
    kind: str = "gbt"


@dataclass
class LinearMember:
    """Binary logistic regression on the raw (unscaled) features (row a10)."""
    name: str
    coef: np.ndarray          # [F] f64
    intercept: float
    kind: str = "linear"


@dataclass
class ModelArrays:
    n_features: int
    members: list             # in estimators_ order == meta-feature column order
    final_coef: np.ndarray    # [n_members] f64
    final_intercept: float
    classes: np.ndarray       # [2] f64
    source: str = ""
    meta: dict = field(default_factory=dict)


# ---------------------------------------------------------------------------------------------------------------------
# uniform attribute access over live estimators and inert stand-ins

def _has(obj, name):
    if isinstance(obj, Inert):
        return isinstance(obj.state, dict) and name in obj.state
    return hasattr(obj, name)


def _get(obj, name):
    if isinstance(obj, Inert):
        return unwrap(obj[name])
    try:
        return getattr(obj, name)
    except AttributeError as exc:  # unfitted estimator
        raise ModelShapeError(f"{type(obj).__name__} has no fitted attribute {name!r}") from exc


def _kind(obj):
    if isinstance(obj, Inert):
        return obj.qualname.rsplit(".", 1)[-1]
    return type(obj).__name__


def _f64(a, shape=None, what=""):
    arr = np.array(a, dtype=np.float64, copy=True)
    if shape is not None and arr.shape != tuple(shape):
        raise ModelShapeError(f"{what}: expected shape {tuple(shape)}, found {arr.shape}")
    if not np.all(np.isfinite(arr)):
        raise ModelShapeError(f"{what}: non-finite parameter")
    return arr


def _scalar(a, what=""):
    arr = np.asarray(a, dtype=np.float64).ravel()
    if arr.size != 1 or not np.isfinite(arr[0]):
        raise ModelShapeError(f"{what}: expected one finite number")
    return float(arr[0])


# ---------------------------------------------------------------------------------------------------------------------

def _walk_svc_pipeline(name, pipe, n_features):
    steps = _get(pipe, "steps")
    if len(steps) != 2:
        raise ModelShapeError("SVC member must be Pipeline(StandardScaler, SVC)")
    scaler, svc = steps[0][1], steps[1][1]
    if _kind(scaler) != "StandardScaler" or _kind(svc) != "SVC":
        raise ModelShapeError("SVC member must be Pipeline(StandardScaler, SVC)")
    with_mean = _get(scaler, "with_mean") if _has(scaler, "with_mean") else True
    with_std = _get(scaler, "with_std") if _has(scaler, "with_std") else True
    mean = _f64(_get(scaler, "mean_"), (n_features,), "scaler mean_") if with_mean else np.zeros(n_features)
    scale = _f64(_get(scaler, "scale_"), (n_features,), "scaler scale_") if with_std else np.ones(n_features)
    if np.any(scale == 0.0):
        raise ModelShapeError("scaler scale_ holds a zero")
    if _get(svc, "kernel") != "rbf":
        raise ModelShapeError("only the RBF kernel of the reference model is implemented")
    if _has(svc, "_sparse") and _get(svc, "_sparse"):
        raise ModelShapeError("sparse-fitted SVC is not supported")
    sv = _f64(_get(svc, "support_vectors_"), None, "support_vectors_")
    if sv.ndim != 2 or sv.shape[1] != n_features or sv.shape[0] < 1:
        raise ModelShapeError("support_vectors_ has the wrong shape")
    n_sv = sv.shape[0]
    n_support = np.asarray(_get(svc, "_n_support"), dtype=np.int64).ravel()
    # the CVE-2020-28975 guard sklearn applies before calling libsvm (svm/_base.py:680-686)
    if n_support.size != 2 or int(n_support.sum()) != n_sv or np.any(n_support < 0):
        raise ModelShapeError("SVC is inconsistent: n_support does not add up to the support vector count")
    dual = _f64(_get(svc, "_dual_coef_"), (1, n_sv), "_dual_coef_")[0]
    prob_a = np.asarray(_get(svc, "_probA"), dtype=np.float64).ravel()
    prob_b = np.asarray(_get(svc, "_probB"), dtype=np.float64).ravel()
    if prob_a.size != 1 or prob_b.size != 1:
        raise ModelShapeError("SVC was not fitted with probability=True")
    gamma = _scalar(_get(svc, "_gamma"), "_gamma")
    if gamma <= 0:
        raise ModelShapeError("_gamma must be positive")
    return SvcMember(name=name, mean=mean, scale=scale, gamma=gamma, sv=sv, dual_coef=dual,
                     intercept=_scalar(_get(svc, "_intercept_"), "_intercept_"),
                     prob_a=float(prob_a[0]), prob_b=float(prob_b[0]), n_support=n_support)


def tree_table_from_state(nodes, values, n_features):
    """``Tree.__getstate__`` record array (7-field 0.23 layout or 8-field 1.x layout) -> TreeTable."""
    nodes = np.asarray(nodes)
    names = nodes.dtype.names or ()
    for need in ("left_child", "right_child", "feature", "threshold"):
        if need not in names:
            raise ModelShapeError(f"tree node records lack {need!r}")
    n = nodes.shape[0]
    left = nodes["left_child"].astype(np.int64)
    right = nodes["right_child"].astype(np.int64)
    feat = nodes["feature"].astype(np.int64)
    thr = nodes["threshold"].astype(np.float64)
    val = np.asarray(values, dtype=np.float64)
    if val.ndim != 3 or val.shape != (n, 1, 1):
        raise ModelShapeError("tree values must have shape [n_nodes,1,1]")
    val = val[:, 0, 0].copy()
    leaf = left == -1
    if n < 1 or np.any((right == -1) != leaf):
        raise ModelShapeError("tree has a half-leaf node")
    inner = ~leaf
    if np.any(left[inner] <= np.nonzero(inner)[0]) or np.any(right[inner] <= np.nonzero(inner)[0]):
        raise ModelShapeError("tree children must come after their parent (no cycles)")
    if np.any(left[inner] >= n) or np.any(right[inner] >= n):
        raise ModelShapeError("tree child index out of range")
    if np.any(feat[inner] < 0) or np.any(feat[inner] >= n_features):
        raise ModelShapeError("tree split feature out of range")
    if not np.all(np.isfinite(thr[inner])) or not np.all(np.isfinite(val)):
        raise ModelShapeError("tree holds a non-finite threshold or value")
    feat = np.where(leaf, -2, feat)
    return TreeTable(left=left.astype(np.int32), right=right.astype(np.int32), feature=feat.astype(np.int32),
                     threshold=np.where(leaf, -2.0, thr), value=val)


def _tree_state(tree_obj):
    if isinstance(tree_obj, Inert):
        st = tree_obj.state
        return unwrap(st["nodes"]), unwrap(st["values"])
    st = tree_obj.__getstate__()  # plain dict of arrays; no model code beyond the accessor
    return st["nodes"], st["values"]


def _walk_gbt(name, gbc, n_features):
    n_classes = int(_get(gbc, "n_classes_"))
    if n_classes != 2:
        raise ModelShapeError("boosting member must be binary")
    ests = np.asarray(_get(gbc, "estimators_"), dtype=object)
    if ests.ndim != 2 or ests.shape[1] != 1:
        raise ModelShapeError("boosting member must have K=1 tree per stage")
    loss = _get(gbc, "loss")
    if loss not in ("deviance", "log_loss"):
        raise ModelShapeError(f"boosting loss {loss!r} is not the binomial deviance")
    init = _get(gbc, "init_")
    if isinstance(init, str):
        if init != "zero":
            raise ModelShapeError("unknown init_")
        init_raw = 0.0
    else:
        if _kind(init) != "DummyClassifier":
            raise ModelShapeError("init_ must be the default prior DummyClassifier")
        prior = _f64(_get(init, "class_prior_"), (2,), "class_prior_")
        eps = np.finfo(np.float64).eps                      # ensemble/_gb.py:125-126
        p1 = float(np.clip(prior[1], eps, 1 - eps))
        init_raw = float(np.log(p1 / (1.0 - p1)))           # _loss/link.py:172-181 (scipy logit)
    lr = _scalar(_get(gbc, "learning_rate"), "learning_rate")
    trees = []
    for stage in ests[:, 0]:
        nodes, values = _tree_state(_get(stage, "tree_"))
        trees.append(tree_table_from_state(nodes, values, n_features))
    return GbtMember(name=name, init_raw=init_raw, learning_rate=lr, trees=trees)


def _walk_logreg(name, lr, n_in):
    coef = _f64(_get(lr, "coef_"), (1, n_in), f"{name} coef_")[0]
    icpt = np.asarray(_get(lr, "intercept_"), dtype=np.float64).ravel()
    if icpt.size != 1:
        raise ModelShapeError("logistic member must be binary")
    return LinearMember(name=name, coef=coef, intercept=_scalar(icpt, "intercept_"))


def _n_features_of(member):
    kind = _kind(member)
    if kind == "Pipeline":
        return int(np.asarray(_get(_get(member, "steps")[0][1], "mean_")).shape[0])
    if kind == "GradientBoostingClassifier":
        for key in ("n_features_in_", "n_features_"):
            if _has(member, key):
                return int(_get(member, key))
    if kind == "LogisticRegression":
        return int(np.asarray(_get(member, "coef_")).shape[1])
    raise ModelShapeError(f"unsupported stack member {kind}")


def arrays_from_stack(stack, source="") -> ModelArrays:
    """Walk a fitted stack (live sklearn object or inert decode) into ``ModelArrays``."""
    if _kind(stack) != "StackingClassifier":
        raise ModelShapeError(f"expected a StackingClassifier, found {_kind(stack)}")
    if _has(stack, "passthrough") and _get(stack, "passthrough"):
        raise ModelShapeError("passthrough=True stacks are not implemented")
    methods = list(_get(stack, "stack_method_"))
    members = list(_get(stack, "estimators_"))
    if len(members) < 1 or len(methods) != len(members) or any(m != "predict_proba" for m in methods):
        raise ModelShapeError("every member must be stacked through predict_proba")
    if _has(stack, "estimators"):
        names = [str(pair[0]) for pair in _get(stack, "estimators") if not isinstance(pair[1], str)]
    else:
        names = []
    if len(names) != len(members):
        names = [f"m{i}" for i in range(len(members))]
    n_features = _n_features_of(members[0])
    out = []
    for name, est in zip(names, members):
        kind = _kind(est)
        if _n_features_of(est) != n_features:
            raise ModelShapeError("members disagree on the number of features")
        if kind == "Pipeline":
            out.append(_walk_svc_pipeline(name, est, n_features))
        elif kind == "GradientBoostingClassifier":
            out.append(_walk_gbt(name, est, n_features))
        elif kind == "LogisticRegression":
            out.append(_walk_logreg(name, est, n_features))
        else:
            raise ModelShapeError(f"unsupported stack member {kind}")
    final = _get(stack, "final_estimator_")
    if _kind(final) != "LogisticRegression":
        raise ModelShapeError("the combiner must be a LogisticRegression")
    fin = _walk_logreg("final", final, len(out))
    classes = _f64(_get(stack, "classes_"), (2,), "classes_")
    version = _get(stack, "_sklearn_version") if _has(stack, "_sklearn_version") else ""
    return ModelArrays(n_features=n_features, members=out, final_coef=fin.coef, final_intercept=fin.intercept,
                       classes=classes, source=source, meta={"sklearn_version": str(version)})


def arrays_from_published_pickle(path) -> ModelArrays:
    """Static decode of ``hf_predict_model.pkl`` (never unpickled with real classes) -> ``ModelArrays``."""
    from .pickle_static import decode_pickle_file
    return arrays_from_stack(decode_pickle_file(path), source=f"static:{path}")
