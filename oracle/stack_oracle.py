"""CPU oracle: numpy float64 restatement of the reference's scoring path.  TEST INFRASTRUCTURE ONLY.

Nothing in the product (``machine-learning-replications_b200/``) imports this module; only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference`` legs may.  It is the checker, never
the thing that is shipped or measured as the GPU path.

The arithmetic of the reference's hot path does not live under /root/reference: predict_hf.py:36 and
train_ensemble_public.py:62 call ``StackingClassifier.predict_proba`` of scikit-learn (un-vendored, unpinned; the
pickle says 0.23.2, this image has 1.9.0).  Every function below cites the scikit-learn 1.9.0 source it restates
(``SK/`` = site-packages/sklearn).  Parity pinning: the reference ships no tests or golden vectors (SURVEY.md section
8c), so the oracle is pinned against (i) scikit-learn 1.9.0 itself, run in this image on estimators rebuilt from the
same arrays (``oracle/sk_rebuild.py``; tests/test_oracle.py), and (ii) the committed fixtures in tests/golden/ that
``oracle/make_golden.py`` generated from scikit-learn.
"""
from __future__ import annotations

import numpy as np
from scipy.special import expit


# --- SVC member -------------------------------------------------------------------------------------------------------

def scaler_transform(m, X):
    """SK/preprocessing/_data.py:1132-1134: ``X -= mean_; X /= scale_`` in float64."""
    return (np.asarray(X, dtype=np.float64) - m.mean) / m.scale


def svc_decision(m, X, chunk=4096):
    """SK/svm/src/libsvm/svm.cpp:2821-2904 (predict_values) with the RBF k_function of svm.cpp:461-515:
    ``f = sum_j coef_j * exp(-gamma * |z - sv_j|^2) - rho`` where ``-rho`` is sklearn's ``_intercept_``."""
    Z = scaler_transform(m, X)
    out = np.empty(Z.shape[0], dtype=np.float64)
    for s in range(0, Z.shape[0], chunk):
        d = Z[s:s + chunk, None, :] - m.sv[None, :, :]
        k = np.exp(-m.gamma * np.einsum("nsf,nsf->ns", d, d))
        out[s:s + chunk] = k @ m.dual_coef + m.intercept
    return out


def sigmoid_predict(f, a, b):
    """svm.cpp:2035-2043: sign-split logistic of ``f*A + B``."""
    t = f * a + b
    e = np.exp(-np.abs(t))
    return np.where(t >= 0, e / (1.0 + e), 1.0 / (1.0 + e))


def coupling_2class(r01):
    """svm.cpp:2046-2107 (multiclass_probability, Wu-Lin-Weng method 2) for k=2, vectorised over rows.

    Returns ``(p, iters)`` with ``p[:, t]`` the coupled probability of libsvm class ``t``.  The early exit
    ``max_error < eps`` (eps = 0.005/k) is per row; rows that have exited are frozen."""
    r01 = np.asarray(r01, dtype=np.float64)
    n = r01.shape[0]
    r10 = 1.0 - r01
    q00, q11 = r10 * r10, r01 * r01         # Q[t][t] = sum_{j != t} r[j][t]^2
    q01 = -r10 * r01                        # Q[0][1] = -r[1][0]*r[0][1]; Q[1][0] = Q[0][1]
    p0 = np.full(n, 0.5)
    p1 = np.full(n, 0.5)
    iters = np.zeros(n, dtype=np.int32)
    active = np.ones(n, dtype=bool)
    eps = 0.005 / 2
    for _ in range(100):
        qp0 = q00 * p0 + q01 * p1
        qp1 = q01 * p0 + q11 * p1
        pqp = p0 * qp0 + p1 * qp1
        err = np.maximum(np.abs(qp0 - pqp), np.abs(qp1 - pqp))
        active &= ~(err < eps)
        if not active.any():
            break
        a = active
        # t = 0
        diff = (-qp0[a] + pqp[a]) / q00[a]
        n0 = p0[a] + diff
        pq = (pqp[a] + diff * (diff * q00[a] + 2 * qp0[a])) / (1 + diff) / (1 + diff)
        nq0 = (qp0[a] + diff * q00[a]) / (1 + diff)
        nq1 = (qp1[a] + diff * q01[a]) / (1 + diff)
        n0 = n0 / (1 + diff)
        n1 = p1[a] / (1 + diff)
        # t = 1
        diff = (-nq1 + pq) / q11[a]
        n1 = n1 + diff
        pq = (pq + diff * (diff * q11[a] + 2 * nq1)) / (1 + diff) / (1 + diff)
        n0 = n0 / (1 + diff)
        n1 = n1 / (1 + diff)
        p0[a], p1[a] = n0, n1
        iters[a] += 1
    return np.stack([p0, p1], axis=1), iters


def svc_predict_proba1(m, X):
    """svm.cpp:2921-2958: Platt value, clamp to [1e-7, 1-1e-7], coupling; returns P(class 1)."""
    r = np.clip(sigmoid_predict(svc_decision(m, X), m.prob_a, m.prob_b), 1e-7, 1 - 1e-7)
    return coupling_2class(r)[0][:, 1]


# --- boosting member --------------------------------------------------------------------------------------------------

def tree_apply(t, X32):
    """SK/ensemble/_gradient_boosting.pyx:25-73 / SK/tree/_tree.pyx:954-996: float32 feature vs float64 threshold,
    ``<=`` goes left.  Returns the sklearn node id of the leaf each row lands in."""
    node = np.zeros(X32.shape[0], dtype=np.int64)
    rows = np.arange(X32.shape[0])
    left, right, feat, thr = (t.left.astype(np.int64), t.right.astype(np.int64), t.feature.astype(np.int64),
                              t.threshold)
    while True:
        inner = left[node] != -1
        if not inner.any():
            return node
        nd = node[inner]
        go_left = X32[rows[inner], feat[nd]].astype(np.float64) <= thr[nd]
        node[inner] = np.where(go_left, left[nd], right[nd])


def gbt_apply(m, X):
    """SK/ensemble/_gb.py:1104-1135 (apply): leaf ids [n, n_trees] (sklearn returns them as float64 [n,T,1])."""
    X32 = np.ascontiguousarray(X, dtype=np.float32)
    return np.stack([tree_apply(t, X32) for t in m.trees], axis=1)


def gbt_raw(m, X):
    """SK/ensemble/_gb.py:981-986 + _gradient_boosting.pyx:164-206: ``raw = init + sum_t lr * value_t[leaf]``,
    float64 accumulation in boosting order."""
    X32 = np.ascontiguousarray(X, dtype=np.float32)
    raw = np.full(X32.shape[0], m.init_raw, dtype=np.float64)
    for t in m.trees:
        raw += m.learning_rate * t.value[tree_apply(t, X32)]
    return raw


# --- linear members and the stack -------------------------------------------------------------------------------------

def linear_decision(m, X):
    """SK/linear_model/_base.py:366-396: ``X @ coef_.T + intercept_`` on the raw features."""
    return np.asarray(X, dtype=np.float64) @ m.coef + m.intercept


def member_proba1(m, X):
    if m.kind == "svc":
        return svc_predict_proba1(m, X)
    if m.kind == "gbt":
        return expit(gbt_raw(m, X))                 # SK/_loss/loss.py:980-998
    if m.kind == "linear":
        return expit(linear_decision(m, X))         # SK/linear_model/_base.py:429-451
    raise ValueError(m.kind)


def transform(model, X):
    """SK/ensemble/_stacking.py:292-300, 87-135: column 1 of every member's predict_proba -> [n, n_members]."""
    return np.stack([member_proba1(m, X) for m in model.members], axis=1)


def decision_function(model, X):
    return transform(model, X) @ model.final_coef + model.final_intercept


def predict_proba(model, X):
    """SK/ensemble/_stacking.py:768-794 -> final LogisticRegression (_base.py:429-451): ``[1-p, p]``."""
    p = expit(decision_function(model, X))
    return np.stack([1.0 - p, p], axis=1)


def predict(model, X):
    """SK/ensemble/_stacking.py:711-766 -> SK/linear_model/_base.py:398-425: strict ``z > 0`` then classes_."""
    return model.classes[(decision_function(model, X) > 0).astype(np.int64)]
