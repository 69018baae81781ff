"""Model export layer: static pickle decoder, attribute walker, flat blob.  CPU only."""
import io
import os
import pickle
import warnings
import zlib

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from conftest import REFERENCE_DIR
from oracle import stack_oracle as so


def test_static_decoder_refuses_foreign_globals(pkg):
    class Evil:
        def __reduce__(self):
            return (os.system, ("echo pwned",))

    data = pickle.dumps({"x": Evil()}, protocol=3)
    with pytest.raises(pkg.export.StaticDecodeError):
        pkg.export.decode_pickle_bytes(data)
    with pytest.raises(pkg.export.StaticDecodeError):
        import collections
        pkg.export.decode_pickle_bytes(pickle.dumps(collections.OrderedDict(a=1), protocol=3))   # not on the allow-list
    with pytest.raises(pkg.export.StaticDecodeError):
        pkg.export.decode_pickle_bytes(pickle.dumps([1, 2, 3], protocol=3))      # top level is not an estimator


@pytest.mark.skipif(not os.path.isdir(REFERENCE_DIR), reason="reference tree not mounted on this machine")
def test_published_pickle_decodes_to_committed_blob(pkg, published):
    arrays = pkg.export.arrays_from_published_pickle(os.path.join(REFERENCE_DIR, "hf_predict_model.pkl"))
    assert pkg.export.write_blob(arrays) == published["blob"]
    svc, gbt, lin = arrays.members
    assert (svc.kind, gbt.kind, lin.kind) == ("svc", "gbt", "linear")
    assert svc.sv.shape == (434, 17) and svc.n_support.tolist() == [321, 113]
    assert abs(svc.gamma - 1 / 17) < 1e-15 and abs(svc.prob_a - (-1.258577321487791)) < 1e-15
    assert len(gbt.trees) == 100 and all(t.left.shape[0] == 3 for t in gbt.trees)
    assert abs(gbt.init_raw - (-1.4003791010016295)) < 1e-12
    assert np.allclose(arrays.final_coef, [1.8372434, 0.41020655, 2.88042418], atol=1e-8)


def test_blob_roundtrip_and_validation(pkg, case):
    blob = case["blob"]
    again = pkg.export.write_blob(pkg.export.read_blob(blob))
    assert again == blob
    bad = bytearray(blob)
    bad[len(bad) // 2] ^= 0x40
    with pytest.raises(pkg.export.BlobError, match="checksum"):
        pkg.export.read_blob(bytes(bad))
    with pytest.raises(pkg.export.BlobError):
        pkg.export.read_blob(blob[:-16])
    with pytest.raises(pkg.export.BlobError):
        pkg.export.read_blob(b"NOTABLOB" + blob[8:])


def test_walker_on_live_sklearn_stack(pkg):
    from sklearn.ensemble import GradientBoostingClassifier, StackingClassifier
    from sklearn.linear_model import LogisticRegression
    from sklearn.pipeline import make_pipeline
    from sklearn.preprocessing import StandardScaler
    from sklearn.svm import SVC

    specs = pkg.synth.schema_specs()
    X = pkg.synth.host_rows(specs, 5, 0, 400).astype(np.float64)
    y = (pkg.synth.hash_u32(5, np.arange(400), 250) < 2 ** 30).astype(float)
    y[X[:, 3] > 0] = 1.0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        clf = StackingClassifier(
            estimators=[("svc", make_pipeline(StandardScaler(), SVC(class_weight="balanced", probability=True,
                                                                    random_state=0))),
                        ("gbc", GradientBoostingClassifier(n_estimators=20, max_depth=3, random_state=0)),
                        ("lg", LogisticRegression(class_weight="balanced", penalty="l1", solver="liblinear"))],
            final_estimator=LogisticRegression(class_weight="balanced")).fit(X, y)
        arrays = pkg.export.arrays_from_stack(clf)
        Xt = pkg.synth.host_rows(specs, 6, 1000, 500).astype(np.float64)
        assert np.abs(clf.predict_proba(Xt) - so.predict_proba(arrays, Xt)).max() < 1e-12
        assert np.array_equal(clf.predict(Xt), so.predict(arrays, Xt))
    # tampered SVC (the CVE-2020-28975 shape check) is refused
    clf.estimators_[0].steps[1][1]._n_support = np.array([1, 1], dtype=np.int32)
    with pytest.raises(pkg.export.ModelShapeError):
        pkg.export.arrays_from_stack(clf)


def test_walker_rejects_other_shapes(pkg):
    from sklearn.linear_model import LogisticRegression
    with pytest.raises(pkg.export.ModelShapeError):
        pkg.export.arrays_from_stack(LogisticRegression())


def test_load_cohort_mat_matches_reference_loader(pkg, tmp_path):
    import scipy.io as sio
    data = np.arange(60, dtype=np.int32).reshape(10, 6)
    path = str(tmp_path / "cohort.mat")
    sio.savemat(path, {"data_tb": data, "clin_var_names": np.array([["a", "b", "c", "d", "e"]], dtype=object)})
    X, Y, names = pkg.load_cohort_mat(path)
    assert X.dtype == np.float64 and Y.dtype == np.float64 and X.shape == (10, 5)
    assert np.array_equal(Y, data[:, -1])
    if os.path.isdir(REFERENCE_DIR):
        import sys
        sys.path.insert(0, REFERENCE_DIR)
        try:
            from load_data_public import load_data
        finally:
            sys.path.remove(REFERENCE_DIR)
        Xr, Yr, _ = load_data(path)
        assert np.array_equal(X, Xr) and np.array_equal(Y, Yr)


def test_patient_vector_order(pkg):
    params = {k: i for i, k in enumerate(pkg.synth.SCHEMA_NAMES)}
    assert pkg.patient_vector(params).tolist() == [list(range(17))]
    with pytest.raises(ValueError):
        pkg.patient_vector({"Gender": 1})


@settings(max_examples=50, deadline=None)
@given(seed=st.integers(0, 2 ** 40), row0=st.integers(0, 2 ** 33), n=st.integers(1, 64))
def test_synth_rows_are_a_pure_function_of_row_index(pkg, seed, row0, n):
    specs = pkg.synth.schema_specs()
    a = pkg.synth.host_rows(specs, seed, row0, n)
    b = pkg.synth.host_rows(specs, seed, 0, 0, rows=np.arange(row0, row0 + n)[::-1])[::-1]
    assert a.dtype == np.float32 and np.array_equal(a, b)
    assert a[:, 6].min() >= 1 and a[:, 6].max() <= 2 and a[:, 15].max() <= 4 and a[:, 13].min() >= 8
