"""The oracle against what scikit-learn 1.9.0 itself returns (committed fixtures + live rebuilt estimators) and the
known answers of SURVEY.md Appendix A.  CPU only."""
import warnings

import numpy as np
import pytest

from oracle import sk_rebuild, stack_oracle as so


def test_oracle_matches_sklearn_golden(case):
    g, m = case["golden"], case["arrays"]
    X = g["X"].astype(np.float64)
    assert np.abs(so.predict_proba(m, X) - g["proba"]).max() < 1e-12
    assert np.array_equal(so.predict(m, X), g["label"])
    assert np.abs(so.transform(m, X) - g["meta"]).max() < 1e-12
    gbt = next(mm for mm in m.members if mm.kind == "gbt")
    assert np.array_equal(so.gbt_apply(gbt, X), g["leaves"].astype(np.int64))


def test_appendix_known_answers(published):
    m = published["arrays"]
    X = published["golden"]["X"][:3].astype(np.float64)
    assert X[0].tolist() == [1, 1, 0, 0, 1, 0, 1, 1, 0, 0, 0, 0, 0, 13, 0, 0, 55]      # predict_hf.py:6-26 defaults
    p1 = so.predict_proba(m, X)[:, 1]
    assert abs(p1[0] - 0.27090030089940764) < 1e-13                                       # prints "27.09 %"
    assert abs(p1[1] - 0.14744957101508452) < 1e-13
    assert abs(p1[2] - 0.8435903076682638) < 1e-13
    assert so.predict(m, X).tolist() == [0.0, 0.0, 1.0]
    meta = so.transform(m, X)
    assert np.allclose(meta[0], [0.08854113301737558, 0.09889406359831056, 0.2763945829171973], atol=1e-13)
    assert so.gbt_apply(m.members[1], X)[0, :6].tolist() == [1, 1, 1, 1, 2, 1]
    assert abs(so.svc_decision(m.members[0], X[:1])[0] - 0.9072594486151935) < 1e-12
    assert abs(so.gbt_raw(m.members[1], X[:1])[0] - (-2.209573614626111)) < 1e-12


def test_oracle_matches_live_sklearn_on_random_rows(case, pkg):
    m = case["arrays"]
    X = pkg.synth.host_rows(case["specs"], 77, 5_000_000, 3000).astype(np.float64)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        st = sk_rebuild.rebuild_stack(m)
        assert np.abs(st.predict_proba(X) - so.predict_proba(m, X)).max() < 1e-12
        assert np.array_equal(st.predict(X), so.predict(m, X))
        gi = next(i for i, mm in enumerate(m.members) if mm.kind == "gbt")
        assert np.array_equal(st.estimators_[gi].apply(X)[:, :, 0], so.gbt_apply(m.members[gi], X))


def test_coupling_plateau_and_edges():
    # svm.cpp:2046-2107: rows whose Platt value lies in [0.495, 0.505] exit before the first sweep -> exactly 0.5
    r = np.array([0.4951, 0.5, 0.5049, 0.4949, 0.5051, 1e-7, 1 - 1e-7, 0.9, 0.1])
    p, it = so.coupling_2class(r)
    assert np.all(p[:3] == 0.5) and np.all(it[:3] == 0)
    assert np.all(it[3:5] >= 1) and np.all(p[3:5, 1] != 0.5)
    assert np.all(np.abs(p.sum(1) - 1) < 1e-12)
    assert it.max() <= 2
    # the fixed point is p = [r, 1-r]; the early stop leaves at most ~0.005 of error
    assert np.abs(p[:, 0] - r).max() < 0.0051


def test_float32_rounding_contract(published):
    # the boosting member compares float32(x) with the float64 threshold; SVC and logistic members see float64 x
    m = published["arrays"]
    x = np.array([[0, 0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 20.5, 0, 0, 38.75]], dtype=np.float64)
    x2 = x.copy()
    x2[0, 13] = np.nextafter(np.float32(20.5), np.float32(30)).astype(np.float64)
    la, lb = so.gbt_apply(m.members[1], x), so.gbt_apply(m.members[1], x2)
    assert (la != lb).any()                      # 20.5 is a published threshold: <= goes left, next float32 goes right
