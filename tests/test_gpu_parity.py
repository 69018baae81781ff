"""Parity of the CUDA path against the oracle and the committed scikit-learn fixtures, through the C ABI.

Bars (BASELINE.json north_star): leaf indices and hard labels bit-exact; probabilities within 1e-6 absolute.
"""
import ctypes as C
import warnings

import numpy as np
import pytest

from oracle import stack_oracle as so

pytestmark = pytest.mark.gpu
PROBA_TOL = 1e-6          # absolute, float32 output vs scikit-learn's float64


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "these tests need a CUDA device"
    return torch


@pytest.fixture(scope="module")
def clf_for(pkg, cuda_lib, torch_cuda):
    cache = {}

    def make(case):
        if case["name"] not in cache:
            cache[case["name"]] = pkg.HFStackB200(case["blob"], device=0)
        return cache[case["name"]]

    return make


def device_scores(clf, X):
    """The device protocol (ingest kernel -> fused / streaming kernel -> fix-up) whatever the row count; host calls
    of up to 4096 rows take the zero-copy float64 path instead."""
    import torch
    n = len(X)
    x = clf.to_device_colmajor(X)
    out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
    lab = torch.empty(n, dtype=torch.uint8, device="cuda")
    clf.score_device(x, n=n, out=out, labels=lab)
    torch.cuda.synchronize()
    return out.cpu().numpy().astype(np.float64), lab.cpu().numpy()


def test_golden_fixture_parity(case, clf_for):
    clf, g = clf_for(case), case["golden"]
    X = g["X"].astype(np.float64)
    proba = clf.predict_proba(X)
    assert proba.dtype == np.float64 and proba.shape == (X.shape[0], 2)
    assert np.abs(proba - g["proba"]).max() <= PROBA_TOL
    assert np.array_equal(proba[:, 0], 1.0 - proba[:, 1])              # p0 is computed from p1, not rounded on its own
    assert np.array_equal(clf.predict(X), g["label"])                  # hard labels bit-exact, edge rows included
    assert np.array_equal(clf.apply(X), g["leaves"].astype(np.int64))  # leaf indices bit-exact
    assert np.abs(clf.transform(X) - g["meta"]).max() <= 2e-6
    assert np.abs(clf.decision_function(X) - so.decision_function(case["arrays"], X)).max() <= 5e-6
    assert np.array_equal(clf.classes_, [0.0, 1.0]) and clf.n_features_in_ == 17
    # the same rows through the device protocol: float32 RBF sum, guard bands and the work-list fix-up at work on the
    # rows the fixture places next to every discontinuity
    dev, dev_labels = device_scores(clf, X)
    assert np.abs(dev - g["proba"]).max() <= PROBA_TOL
    assert np.array_equal(clf.classes_[dev_labels], g["label"])


def test_predict_hf_default_patient(published, clf_for, pkg):
    # what predict_hf.py:29-40 does, with the facade in place of the unpickled model
    params = dict(zip(pkg.synth.SCHEMA_NAMES, [1, 1, 0, 0, 1, 0, 1, 1, 0, 0, 0, 0, 0, 13, 0, 0, 55]))
    score = clf_for(published).predict_proba(pkg.patient_vector(params))
    assert abs(score[0, 1] - 0.27090030089940764) <= PROBA_TOL
    assert "{0:.2f}".format(100 * score[0, 1]) == "27.09"


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 31, 32, 33, 127, 1023, 1024, 1025, 4095, 4096, 4097, 4099])
def test_ragged_row_counts(published, clf_for, pkg, n):
    X = pkg.synth.host_rows(published["specs"], 11, 777, n).astype(np.float64)
    want, want_labels = so.predict_proba(published["arrays"], X), so.predict(published["arrays"], X)
    got = clf_for(published).predict_proba(X)                    # host protocol
    assert np.abs(got - want).max() <= PROBA_TOL
    assert np.array_equal(clf_for(published).predict(X), want_labels)
    dev, dev_labels = device_scores(clf_for(published), X)       # device protocol: the fused kernel at this size
    assert np.abs(dev - want).max() <= PROBA_TOL
    assert np.array_equal(published["arrays"].classes[dev_labels], want_labels)


def test_input_layouts_and_dtypes(published, clf_for, pkg):
    clf = clf_for(published)
    X = pkg.synth.host_rows(published["specs"], 3, 0, 200)
    base = clf.predict_proba(X.astype(np.float64))
    assert np.array_equal(clf.predict_proba(X), base)                                  # float32
    assert np.array_equal(clf.predict_proba(np.asfortranarray(X.astype(np.float64))), base)
    assert np.array_equal(clf.predict_proba(X.astype(np.int64)), base)                 # predict_hf.py passes ints
    assert np.array_equal(clf.predict_proba(X.tolist()), base)


def test_error_behaviour(published, clf_for):
    clf = clf_for(published)
    with pytest.raises(ValueError):
        clf.predict_proba(np.zeros(17))
    with pytest.raises(ValueError):
        clf.predict_proba(np.zeros((3, 16)))
    with pytest.raises(ValueError):
        clf.predict_proba(np.zeros((0, 17)))
    bad = np.zeros((5, 17))
    bad[2, 4] = np.nan
    with pytest.raises(ValueError):
        clf.predict_proba(bad)
    big = np.zeros((5, 17))
    big[0, 0] = 1e300                           # finite in float64, infinite once cast to float32 (GBC cast)
    with pytest.raises(ValueError):
        clf.predict_proba(big)


def test_device_status_flags_nonfinite(published, clf_for, torch_cuda):
    torch = torch_cuda
    clf = clf_for(published)
    x = torch.zeros((17, 64), dtype=torch.float32, device="cuda")
    x[6] = 1.0
    x[3, 10] = float("inf")
    st = clf.new_status()
    out = clf.predict_proba_device(x, n=64, status=st)
    torch.cuda.synchronize()
    assert int(st[0].item()) & 1
    assert torch.isnan(out[10]).all() and torch.isfinite(out[11]).all()


def test_misaligned_or_short_matrix_is_refused(published, clf_for, pkg, torch_cuda):
    torch = torch_cuda
    clf = clf_for(published)
    lib = pkg.capi.load_library()
    x = torch.zeros((17, 66), dtype=torch.float32, device="cuda")
    out = torch.zeros((64, 2), dtype=torch.float32, device="cuda")
    st = clf.new_status()
    rc = lib.hfe_predict_proba(clf._handle, x.data_ptr(), 64, 66, out.data_ptr(), st.data_ptr(), None)
    assert rc == 1 and b"aligned" in lib.hfe_last_error_string()
    rc = lib.hfe_predict_proba(clf._handle, x.data_ptr(), 64, 32, out.data_ptr(), st.data_ptr(), None)
    assert rc == 1
    rc = lib.hfe_predict_proba(clf._handle, x.data_ptr(), 0, 64, out.data_ptr(), st.data_ptr(), None)
    assert rc == 0                                                  # empty batch is a no-op at the C level


def test_synth_rows_bit_identical_host_device(case, pkg, torch_cuda):
    x = pkg.synth_rows_device(case["specs"], 2020, 999_999_000, 5000)
    host = pkg.synth.host_rows(case["specs"], 2020, 999_999_000, 5000)
    assert np.array_equal(x[:, :5000].T.cpu().numpy(), host)


def test_large_batch_properties(case, clf_for, pkg, torch_cuda):
    """10 M rows (BASELINE config 2): size-independent properties + oracle on a strided sample."""
    torch = torch_cuda
    clf = clf_for(case)
    n = 10_000_000
    x = pkg.synth_rows_device(case["specs"], 2020, 0, n)
    st = clf.new_status()
    out = clf.predict_proba_device(x, n=n, status=st)
    torch.cuda.synchronize()
    assert int(st[0].item()) == 0
    assert int(st[1].item()) < n // 2000                               # float64 re-evaluations stay rare
    assert bool(((out[:, 0] + out[:, 1]) - 1.0).abs().max() <= 1.2e-7)
    assert bool((out >= 0).all()) and bool((out <= 1).all())
    # shard invariance: scoring two halves separately is byte-identical to scoring the whole
    half = n // 2
    a = clf.predict_proba_device(x[:, :half], n=half)
    b = clf.predict_proba_device(x[:, half:], n=n - half)
    torch.cuda.synchronize()
    assert torch.equal(torch.cat([a, b]), out)
    # the float32 fast path agrees with the all-float64 path on every row of a 2 M slice
    m = 2_000_000
    o64 = clf.predict_proba_f64_device(x, n=m)
    torch.cuda.synchronize()
    assert float((out[:m].double() - o64).abs().max().item()) <= 5e-7
    # oracle on a strided sample spread over the whole batch
    idx = np.arange(0, n, n // 20000)[:20000]
    Xs = x[:, torch.from_numpy(idx).cuda()].T.cpu().numpy().astype(np.float64)
    want = so.predict_proba(case["arrays"], Xs)
    got = out[torch.from_numpy(idx).cuda()].cpu().numpy().astype(np.float64)
    assert np.abs(got - want).max() <= PROBA_TOL
    lab = torch.empty(n, dtype=torch.uint8, device="cuda")
    clf.score_device(x, n=n, labels=lab)
    torch.cuda.synchronize()
    assert np.array_equal(lab[torch.from_numpy(idx).cuda()].cpu().numpy(),
                          (so.decision_function(case["arrays"], Xs) > 0).astype(np.uint8))
    # exhaustive: hard labels of the fast pass == labels of the all-float64 pass on EVERY row (SK/linear_model/
    # _base.py:398-425, z > 0), and the margin the label guard really has: the smallest |z| among rows the fast pass
    # did not hand to the float64 fix-up against the largest |z_fast - z_f64| seen on such rows
    dec = torch.empty(n, dtype=torch.float64, device="cuda")
    st2 = clf.new_status()
    clf.score_device(x, n=n, labels=lab, decision=dec, status=st2)
    lab64 = torch.empty(n, dtype=torch.uint8, device="cuda")
    dec64 = torch.empty(n, dtype=torch.float64, device="cuda")
    clf.score_f64_device(x, n=n, labels=lab64, decision=dec64)
    torch.cuda.synchronize()
    assert torch.equal(lab, lab64)
    listed = st2[4:4 + int(st2[2].item())].long()
    assert int(st2[2].item()) <= pkg.capi.WORKLIST_CAPACITY
    unflagged = torch.ones(n, dtype=torch.bool, device="cuda")
    unflagged[listed] = False
    assert float((dec[listed] - dec64[listed]).abs().max().item()) == 0.0        # fix-up rows carry the float64 value
    worst_err = float((dec - dec64).abs()[unflagged].max().item())
    closest = float(dec64.abs()[unflagged].min().item())
    print(f"[{case['name']}] label guard margin: min |z| unflagged {closest:.3e} / max |z32 - z64| {worst_err:.3e} "
          f"= {closest / worst_err:.1f}x; {int(st2[2].item())} rows re-evaluated")
    assert closest > 2.0 * worst_err


def test_solver_edge_rows_take_the_float64_path(published, clf_for, pkg, torch_cuda):
    """Rows engineered to sit on the coupling solver's plateau edge must still match scikit-learn exactly in branch."""
    g = published["golden"]
    X = g["X"][int(g["n_appendix"]) + 1024:].astype(np.float64)        # the edge rows make_golden picked
    clf = clf_for(published)
    want = g["proba"][int(g["n_appendix"]) + 1024:]
    got, _ = device_scores(clf, X)                                      # fused kernel + work-list fix-up
    assert np.abs(got - want).max() <= PROBA_TOL
    assert np.abs(clf.predict_proba(X) - want).max() <= PROBA_TOL       # host protocol


def test_pinned_host_pipeline(case, clf_for, pkg, torch_cuda):
    torch = torch_cuda
    clf = clf_for(case)
    n = 300_001
    X = pkg.synth.host_rows(case["specs"], 9, 0, n)
    xh = torch.from_numpy(X).pin_memory()
    got = clf.predict_proba_pinned(xh, chunk_rows=1 << 16).numpy()
    x = clf.to_device_colmajor(X)
    ref = clf.predict_proba_device(x, n=n).cpu().numpy()
    assert np.array_equal(got, ref)
    got64 = clf.predict_proba_pinned(torch.from_numpy(X.astype(np.float64)).pin_memory(), chunk_rows=1 << 17).numpy()
    assert np.array_equal(got64, ref)
    bad = X.copy()
    bad[123456, 5] = np.nan
    with pytest.raises(ValueError):
        clf.predict_proba_pinned(torch.from_numpy(bad).pin_memory())


def test_large_host_batches_take_the_chunk_pipeline(case, clf_for, pkg, torch_cuda):
    """train_ensemble_public.py:62 on a big cohort: a pageable float64 numpy matrix through hfe_predict_host's chunk
    pipeline (host threads narrow to float32 into pinned slots, three chunks in flight) gives the bytes the device
    protocol gives, for every output kind, at sizes around the chunk boundaries."""
    torch = torch_cuda
    clf = clf_for(case)
    n = (1 << 18) * 2 + 4099                                       # two full chunks and a ragged third
    X = pkg.synth.host_rows(case["specs"], 17, 0, n).astype(np.float64)
    x = clf.to_device_colmajor(X)
    out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
    lab = torch.empty(n, dtype=torch.uint8, device="cuda")
    dec = torch.empty(n, dtype=torch.float64, device="cuda")
    meta = torch.empty((n, 3), dtype=torch.float64, device="cuda")
    clf.score_device(x, n=n, out=out, labels=lab, decision=dec, meta=meta)
    torch.cuda.synchronize()
    proba = clf.predict_proba(X)
    assert proba.dtype == np.float64 and proba.shape == (n, 2)
    assert np.array_equal(proba[:, 1], out[:, 1].cpu().numpy().astype(np.float64))
    assert np.array_equal(proba[:, 0], 1.0 - proba[:, 1])
    assert np.array_equal(clf.predict(X), clf.classes_[lab.cpu().numpy()])
    assert np.array_equal(clf.decision_function(X), dec.cpu().numpy())
    assert np.array_equal(clf.transform(X), meta.cpu().numpy())
    for m in ((1 << 16), (1 << 16) + 1, (1 << 18), (1 << 18) + 1):
        assert np.array_equal(clf.predict_proba(X[:m]), proba[:m]), m
    idx = np.arange(0, n, 97)
    assert np.abs(proba[idx] - so.predict_proba(case["arrays"], X[idx])).max() <= PROBA_TOL
    bad = X[:200_000].copy()
    bad[150_000, 7] = np.inf
    with pytest.raises(ValueError):
        clf.predict_proba(bad)
    assert np.array_equal(clf.predict_proba(X[:70_000]), proba[:70_000])      # the pipeline is reusable after an error
    # wire format: integer cohorts (like X) cross PCIe as bytes; the first chunk holding anything that is not an integer
    # in [0, 255] -- here in the second chunk, after the first went as bytes -- switches the call to float32.  Same bits.
    assert float(np.abs(X - np.rint(X)).max()) == 0.0 and X.min() >= 0 and X.max() <= 255
    for value in (0.5, -1.0, 256.0, 1e9, 254.999999999, -0.0):
        X2 = X.copy()
        X2[(1 << 18) + 12_345, 3] = value
        X2[n - 1, 16] = value
        want, _ = device_scores(clf, X2)
        assert np.array_equal(clf.predict_proba(X2)[:, 1], want[:, 1]), value
    X2[(1 << 18) + 5, 0] = np.nan
    with pytest.raises(ValueError):
        clf.predict_proba(X2)


def test_permuted_column_pass_and_score_count(published, clf_for, pkg, torch_cuda):
    """One permutation-importance pass: column read through an index array == scoring the materialised shuffle."""
    torch = torch_cuda
    clf = clf_for(published)
    n = 100_003
    X = pkg.synth.host_rows(published["specs"], 4, 0, n)
    rng = np.random.RandomState(0)
    perm = rng.permutation(n).astype(np.int32)
    col = 13
    Xp = X.copy()
    Xp[:, col] = X[perm, col]
    x = clf.to_device_colmajor(X)
    xp = clf.to_device_colmajor(Xp)
    y = (so.decision_function(published["arrays"], X.astype(np.float64)) > 0).astype(np.uint8)
    y_dev = torch.from_numpy(y).cuda()
    out_a = torch.empty((n, 2), dtype=torch.float32, device="cuda")
    out_b = torch.empty_like(out_a)
    lab = torch.empty(n, dtype=torch.uint8, device="cuda")
    score = torch.zeros(1, dtype=torch.int64, device="cuda")
    clf.score_device(x, n=n, perm_col=col, perm_idx=torch.from_numpy(perm).cuda(), out=out_a, labels=lab, y=y_dev,
                     score=score)
    clf.score_device(xp, n=n, out=out_b)
    torch.cuda.synchronize()
    assert torch.equal(out_a, out_b)
    assert int(score.item()) == int((lab.cpu().numpy() == y).sum())
    base = torch.zeros(1, dtype=torch.int64, device="cuda")
    clf.score_device(x, n=n, y=y_dev, score=base)
    torch.cuda.synchronize()
    assert int(base.item()) == n                                         # unpermuted labels reproduce y exactly


def test_model_info_and_members(case, clf_for):
    info = clf_for(case).info
    assert info.n_features == 17 and info.n_members == 3
    assert list(info.member_kind)[:3] == [1, 2, 3]
    assert info.n_trees == 100 and info.tree_mode == 1 and 1 <= info.n_merged_splits <= 100
    assert info.n_support_vectors == case["arrays"].members[0].sv.shape[0]


def test_general_tree_models(pkg, clf_for, published, torch_cuda):
    """Depth>1 boosted trees (general node tables) with and without the SVC member, against the oracle."""
    import copy
    rng = np.random.default_rng(5)
    ex = pkg.export

    def random_tree(depth, leaves_target):
        left, right, feat, thr, val = [-1], [-1], [-2], [-2.0], [rng.normal() * 0.3]
        frontier = [(0, 0)]
        n_leaves = 1
        while frontier and n_leaves < leaves_target:
            i, d = frontier.pop(rng.integers(len(frontier)))
            if d >= depth:
                continue
            f = int(rng.integers(17))
            t = {13: rng.integers(10, 30) + 0.5, 16: rng.integers(45, 75) + 0.25, 15: rng.integers(0, 4) + 0.5,
                 6: 1.5}.get(f, 0.5)
            left[i], right[i], feat[i], thr[i] = len(left), len(left) + 1, f, float(t)
            for _ in range(2):
                left.append(-1); right.append(-1); feat.append(-2); thr.append(-2.0); val.append(rng.normal() * 0.3)
            frontier += [(left[i], d + 1), (right[i], d + 1)]
            n_leaves += 1
        return ex.TreeTable(left=np.array(left, np.int32), right=np.array(right, np.int32),
                            feature=np.array(feat, np.int32), threshold=np.array(thr), value=np.array(val))

    base = published["arrays"]
    gbt = ex.GbtMember(name="gbc", init_raw=-1.2, learning_rate=0.05,
                       trees=[random_tree(8, 40) for _ in range(60)])
    with_svc = copy.deepcopy(base)
    with_svc.members[1] = gbt
    no_svc = ex.ModelArrays(n_features=17, members=[gbt, base.members[2]], final_coef=np.array([1.3, 2.1]),
                            final_intercept=-1.1, classes=base.classes)
    X = pkg.synth.host_rows(published["specs"], 21, 0, 20_000).astype(np.float64)
    for arrays in (with_svc, no_svc):
        clf = pkg.HFStackB200.from_arrays(arrays)
        assert clf.info.tree_mode == 2
        assert np.abs(clf.predict_proba(X) - so.predict_proba(arrays, X)).max() <= PROBA_TOL
        assert np.array_equal(clf.predict(X), so.predict(arrays, X))
        g = next(m for m in arrays.members if m.kind == "gbt")
        assert np.array_equal(clf.apply(X[:3000]), so.gbt_apply(g, X[:3000]))
        clf.close()


def test_live_sklearn_object_drop_in(pkg, torch_cuda):
    """train_ensemble_public.py:61-63 with the facade swapped in for clf after fit."""
    from sklearn.ensemble import GradientBoostingClassifier, StackingClassifier
    from sklearn.linear_model import LogisticRegression
    from sklearn.pipeline import make_pipeline
    from sklearn.preprocessing import StandardScaler
    from sklearn.svm import SVC

    specs = pkg.synth.schema_specs()
    X = pkg.synth.host_rows(specs, 5, 0, 500).astype(np.float64)
    y = ((X[:, 3] + X[:, 0] + (X[:, 13] > 20)) >= 2).astype(float)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        clf = StackingClassifier(
            estimators=[("svc", make_pipeline(StandardScaler(), SVC(class_weight="balanced", probability=True,
                                                                    random_state=2020))),
                        ("gbc", GradientBoostingClassifier(n_estimators=100, max_depth=1, random_state=2020)),
                        ("lg", LogisticRegression(class_weight="balanced", penalty="l1", solver="liblinear"))],
            final_estimator=LogisticRegression(class_weight="balanced")).fit(X, y)
        Xs = pkg.synth.host_rows(specs, 6, 10_000, 5000).astype(np.float64)
        fast = pkg.HFStackB200.from_sklearn(clf)
        y_pred = fast.predict_proba(Xs)
        assert np.abs(y_pred - clf.predict_proba(Xs)).max() <= PROBA_TOL
        assert np.array_equal(fast.predict(Xs), clf.predict(Xs))
        gi = 1
        assert np.array_equal(fast.apply(Xs), clf.estimators_[gi].apply(Xs)[:, :, 0].astype(np.int64))


def test_permutation_importance_matches_sklearn(pkg, published, clf_for, torch_cuda):
    """BASELINE config 5 at a size scikit-learn finishes in seconds: same seeds -> same shuffles -> same importances."""
    from sklearn.inspection import permutation_importance as sk_pi

    from oracle import sk_rebuild

    n = 3000
    X = pkg.synth.host_rows(published["specs"], 8, 0, n).astype(np.float64)
    rng = np.random.default_rng(3)
    y = (so.predict_proba(published["arrays"], X)[:, 1] > rng.random(n)).astype(np.float64)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        ref = sk_pi(sk_rebuild.rebuild_stack(published["arrays"]), X, y, n_repeats=3, random_state=7)
    got = pkg.inspection.permutation_importance(clf_for(published), X, y, n_repeats=3, random_state=7)
    assert np.array_equal(got.importances.shape, ref.importances.shape)
    assert np.abs(got.importances - ref.importances).max() < 1e-12        # accuracies are integer counts / n
    assert np.abs(got.importances_mean - ref.importances_mean).max() < 1e-12
    # shuffles drawn on the device: another random stream, the same procedure -- reproducible, statistically the same
    n = 60_000
    X = pkg.synth.host_rows(published["specs"], 9, 0, n).astype(np.float64)
    y = (so.predict_proba(published["arrays"], X)[:, 1] > rng.random(n)).astype(np.float64)
    clf = clf_for(published)
    host = pkg.inspection.permutation_importance(clf, X, y, n_repeats=5, random_state=11)
    dev1 = pkg.inspection.permutation_importance(clf, X, y, n_repeats=5, random_state=11, shuffle="device")
    dev2 = pkg.inspection.permutation_importance(clf, X, y, n_repeats=5, random_state=11, shuffle="device")
    assert np.array_equal(dev1.importances, dev2.importances)
    assert dev1.baseline_score == host.baseline_score
    spread = 4 * np.sqrt(0.25 / n) + 3 * host.importances_std.max()
    assert np.abs(dev1.importances_mean - host.importances_mean).max() < spread
    with pytest.raises(ValueError):
        pkg.inspection.permutation_importance(clf, X, y, shuffle="numpy")


def test_device_metrics_match_sklearn(pkg, published, clf_for, torch_cuda):
    """train_ensemble_public.py:63-64,67,79: classification_report numbers exact, ROC / PR areas to the bin width."""
    from sklearn import metrics as skm
    torch = torch_cuda
    clf = clf_for(published)
    n = 200_000
    X = pkg.synth.host_rows(published["specs"], 12, 0, n)
    x = clf.to_device_colmajor(X)
    proba = clf.predict_proba_device(x, n=n)
    p1 = proba[:, 1].cpu().numpy()
    rng = np.random.default_rng(1)
    y = (p1.astype(np.float64) > rng.random(n)).astype(np.uint8)
    conf, hist = pkg.metrics.device_counts(proba, torch.from_numpy(y).cuda(), threshold=0.5)
    yy = p1 > 0.5                                                   # train:63
    assert np.array_equal(conf, skm.confusion_matrix(y, yy))
    want = skm.classification_report(y, yy, output_dict=True)      # train:64
    got = pkg.metrics.binary_report(conf)
    for key in ("0", "1", "macro avg", "weighted avg"):
        for k2 in ("precision", "recall", "f1-score", "support"):
            assert abs(got[key][k2] - want[key][k2]) < 1e-12
    assert abs(got["accuracy"] - want["accuracy"]) < 1e-12
    assert int(hist.sum()) == n
    assert abs(pkg.metrics.roc_auc(hist) - skm.roc_auc_score(y, p1)) < 2e-5
    assert abs(pkg.metrics.average_precision(hist) - skm.average_precision_score(y, p1)) < 1e-4


def test_work_list_overflow_falls_back_to_float64(pkg, published, clf_for, torch_cuda):
    """More flagged rows than the per-call work list holds: the status word says so and the facade reruns in float64."""
    torch = torch_cuda
    clf = clf_for(published)
    g = published["golden"]
    edge = g["X"][int(g["n_appendix"]) + 1024:].astype(np.float64)          # rows next to a discontinuity
    x = clf.to_device_colmajor(edge)
    st = clf.new_status()
    clf.predict_proba_device(x, n=edge.shape[0], status=st)
    torch.cuda.synchronize()
    flagged = edge[np.sort(st[4:4 + int(st[2])].cpu().numpy())]
    assert flagged.shape[0] >= 1
    reps = pkg.capi.WORKLIST_CAPACITY // flagged.shape[0] + 2
    big = np.tile(flagged, (reps, 1))                                         # > 65532 rows, every one flagged
    xb = clf.to_device_colmajor(big)
    clf.predict_proba_device(xb, n=big.shape[0], status=st)
    torch.cuda.synchronize()
    assert int(st[0].item()) & pkg.capi.WORKLIST_OVERFLOW
    with pytest.raises(pkg.capi.HfeError, match="work list overflow"):        # the facade surfaces it on request ...
        clf.check()
    clf.predict_proba_device(xb, n=big.shape[0], status=st)
    torch.cuda.synchronize()
    with pytest.warns(RuntimeWarning, match="overflowed its work list"):      # ... and on the next device call
        clf.predict_proba_device(x, n=edge.shape[0], status=st)
    torch.cuda.synchronize()
    want = np.tile(so.predict_proba(published["arrays"], flagged), (reps, 1))
    assert np.abs(clf.predict_proba(big) - want).max() <= PROBA_TOL           # host path: automatic float64 rerun
    got = clf.predict_proba_pinned(torch.from_numpy(big.astype(np.float32)).pin_memory()).numpy()
    assert np.abs(got - want).max() <= PROBA_TOL
    assert np.array_equal(clf.predict(big), np.tile(so.predict(published["arrays"], flagged), reps))


def test_large_support_vector_tables(pkg, published, torch_cuda):
    """A support-vector table beyond 48 KB of shared memory (opt-in dynamic smem), and the documented size limit."""
    import copy
    rng = np.random.default_rng(11)
    base = published["arrays"]

    def with_svs(n_sv):
        arrays = copy.deepcopy(base)
        svc = arrays.members[0]
        pick = rng.integers(0, svc.sv.shape[0], n_sv)
        svc.sv = svc.sv[pick] + rng.normal(0, 0.05, (n_sv, 17))
        svc.dual_coef = svc.dual_coef[pick] * (434.0 / n_sv)
        n0 = int((svc.dual_coef <= 0).sum())
        order = np.argsort(svc.dual_coef > 0, kind="stable")
        svc.sv, svc.dual_coef = svc.sv[order], svc.dual_coef[order]
        svc.n_support = np.array([n0, n_sv - n0])
        return arrays

    arrays = with_svs(1203)
    clf = pkg.HFStackB200.from_arrays(arrays)
    X = pkg.synth.host_rows(published["specs"], 13, 0, 6000).astype(np.float64)
    assert np.abs(clf.predict_proba(X) - so.predict_proba(arrays, X)).max() <= PROBA_TOL
    assert np.array_equal(clf.predict(X), so.predict(arrays, X))
    clf.close()
    too_big = pkg.HFStackB200.from_arrays(with_svs(1600))
    with pytest.raises(pkg.capi.HfeError, match="support-vector table"):
        too_big.predict_proba(X[:5000])
    # up to 4096 host rows never touch the shared-memory table (zero-copy float64 path): scored even for this model
    big = with_svs(1600)
    few = pkg.HFStackB200.from_arrays(big).predict_proba(X[:8])
    assert np.abs(few - so.predict_proba(big, X[:8])).max() <= 1e-7      # float64 path, float32 output rounding


def test_prefix_shared_support_vector_layout(pkg, published, retrained, torch_cuda):
    """The prefix-shared table layout (shared partial dot products over the re-ordered leading features): in use for
    both fixture models, falls back to the flat layout when support vectors share nothing, and stays within the
    parity bars on degenerate tables (one vector; every vector sharing one prefix; all-binary vectors with many
    duplicates of a prefix) -- which also exercises the internal feature re-ordering of every member."""
    import copy
    rng = np.random.default_rng(23)
    for fx in (published, retrained):
        clf = pkg.HFStackB200.from_arrays(fx["arrays"])
        n_sv = int(clf.info.n_support_vectors)
        flat = (n_sv + 3) // 4 * 4 * 18
        assert 0 < clf.info.rbf_fma_per_row < 0.75 * flat
        clf.close()

    base = published["arrays"]
    X = pkg.synth.host_rows(published["specs"], 29, 0, 5000).astype(np.float64)

    def variant(make_sv, n_sv):
        arrays = copy.deepcopy(base)
        svc = arrays.members[0]
        pick = rng.integers(0, svc.sv.shape[0], n_sv)
        svc.sv = make_sv(svc.sv[pick].copy())
        svc.dual_coef = svc.dual_coef[pick] * min(434.0 / n_sv, 4.0)
        order = np.argsort(svc.dual_coef > 0, kind="stable")
        svc.sv, svc.dual_coef = svc.sv[order], svc.dual_coef[order]
        n0 = int((svc.dual_coef <= 0).sum())
        svc.n_support = np.array([n0, n_sv - n0])
        return arrays

    def continuous(sv):                      # nothing shared: flat layout
        return sv + rng.normal(0, 0.05, sv.shape)

    def one_prefix(sv):                      # every vector agrees on 13 columns: one class, split into chunks
        sv[:, :13] = sv[0, :13]
        sv[:, 13:] += rng.normal(0, 0.3, sv[:, 13:].shape)
        return sv

    def all_binary(sv):                      # few distinct rows, many exact duplicates
        lo, hi = base.members[0].sv.min(axis=0), base.members[0].sv.max(axis=0)
        bits = rng.integers(0, 2, (6, 17))
        return np.where(bits[rng.integers(0, 6, len(sv))] == 1, hi, lo)

    def mixture(n_binary, n_levels):
        """n_binary two-valued columns (random positions), the rest with n_levels distinct values."""
        def make(sv):
            lo, hi = base.members[0].sv.min(axis=0), base.members[0].sv.max(axis=0)
            cols = rng.permutation(17)
            out = np.empty_like(sv)
            for k, c in enumerate(cols):
                levels = 2 if k < n_binary else n_levels
                grid = np.linspace(lo[c], hi[c], levels)
                out[:, c] = grid[rng.integers(0, levels, len(sv))]
            return out
        return make

    cases = [(continuous, 200, False), (one_prefix, 57, True), (all_binary, 131, True), (lambda sv: sv, 1, None),
             (lambda sv: sv, 3, None), (mixture(17, 2), 300, None), (mixture(12, 5), 900, None),
             (mixture(9, 30), 433, None), (mixture(5, 3), 64, None), (mixture(14, 200), 7, None),
             (mixture(10, 4), 1500, None)]
    for make_sv, n_sv, expect_shared in cases:
        arrays = variant(make_sv, n_sv)
        clf = pkg.HFStackB200.from_arrays(arrays)
        flat = (n_sv + 3) // 4 * 4 * 18
        if expect_shared is True:
            assert clf.info.rbf_fma_per_row < flat
        elif expect_shared is False:
            assert clf.info.rbf_fma_per_row == flat
        assert np.abs(clf.predict_proba(X) - so.predict_proba(arrays, X)).max() <= PROBA_TOL
        assert np.array_equal(clf.predict(X), so.predict(arrays, X))
        assert np.abs(clf.transform(X) - so.transform(arrays, X)).max() <= PROBA_TOL
        gbt = [m for m in arrays.members if m.kind == "gbt"][0]
        assert np.array_equal(clf.apply(X[:500]), so.gbt_apply(gbt, X[:500]))
        clf.close()


def test_tensor_core_engine_against_simt_engine_and_oracle(case, pkg, torch_cuda):
    """The RBF sum on the tensor cores (tcgen05.mma, TMEM accumulators; hfe_mma.cuh) and on the SIMT FFMA2 loop meet the
    same bars: probabilities <= 1e-6 against the oracle, labels exact, the two engines within float32 noise of each
    other -- on schema rows, on values one bfloat16 cannot carry (extension block), with a permuted column, NaN rows,
    rows far outside the training range (exponent bound -> float64 fix-up), every optional output, ragged sizes."""
    torch = torch_cuda
    arrays = case["arrays"]
    clf = pkg.HFStackB200.from_arrays(arrays)
    assert clf.info.rbf_engine == 2                                   # automatic choice for the fixture models
    n = 70_001
    X = pkg.synth.host_rows(case["specs"], 51, 0, n).astype(np.float64)
    rng = np.random.default_rng(4)
    Xf = X + rng.normal(0, 0.41, X.shape) * (rng.random(X.shape) < 0.25)          # fractional values: extension block
    Xo = X.copy()
    Xo[::97] *= 9.0                                                                # outliers: exponent bound
    for name, M in (("schema", X), ("fractional", Xf), ("outliers", Xo)):
        want = so.predict_proba(arrays, M)
        want_lab = (so.decision_function(arrays, M) > 0).astype(np.uint8)
        want_meta = so.transform(arrays, M)
        x = clf.to_device_colmajor(M)
        got = {}
        for eng in ("simt", "tensor"):
            clf.set_rbf_engine(eng)
            assert clf.info.rbf_engine == (1 if eng == "simt" else 2)
            out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
            lab = torch.empty(n, dtype=torch.uint8, device="cuda")
            dec = torch.empty(n, dtype=torch.float64, device="cuda")
            meta = torch.empty((n, 3), dtype=torch.float64, device="cuda")
            st = clf.score_device(x, n=n, out=out, labels=lab, decision=dec, meta=meta)
            torch.cuda.synchronize()
            assert int(st[0].item()) == 0, (name, eng)
            got[eng] = out.cpu().numpy().astype(np.float64)
            assert np.abs(got[eng] - want).max() <= PROBA_TOL, (name, eng)
            assert np.array_equal(lab.cpu().numpy(), want_lab), (name, eng)
            assert np.abs(meta.cpu().numpy() - want_meta).max() <= 2e-6, (name, eng)
            assert np.abs(dec.cpu().numpy() - so.decision_function(arrays, M)).max() <= 5e-6, (name, eng)
            for m in (1, 127, 128, 129, 4099):                           # ragged tiles
                o = clf.predict_proba_device(x[:, :((m + 3) & ~3)].contiguous(), n=m)
                torch.cuda.synchronize()
                assert np.abs(o.cpu().numpy() - want[:m]).max() <= PROBA_TOL, (name, eng, m)
        assert np.abs(got["simt"] - got["tensor"]).max() <= 6e-7, name
    clf.set_rbf_engine("tensor")
    # a permuted column and the on-device accuracy count
    perm = np.random.RandomState(5).permutation(n).astype(np.int32)
    col = 13
    Xp = X.copy()
    Xp[:, col] = X[perm, col]
    x, xp = clf.to_device_colmajor(X), clf.to_device_colmajor(Xp)
    y = (so.decision_function(arrays, Xp) > 0).astype(np.uint8)
    out_a = torch.empty((n, 2), dtype=torch.float32, device="cuda")
    out_b = torch.empty_like(out_a)
    score = torch.zeros(1, dtype=torch.int64, device="cuda")
    clf.score_device(x, n=n, perm_col=col, perm_idx=torch.from_numpy(perm).cuda(), out=out_a, y=torch.from_numpy(y).cuda(),
                     score=score)
    clf.score_device(xp, n=n, out=out_b)
    torch.cuda.synchronize()
    assert torch.equal(out_a, out_b) and int(score.item()) == n
    # NaN / Inf rows are flagged and come out as NaN, their neighbours are untouched
    bad = X[:1000].copy()
    bad[77, 5] = np.nan
    bad[500, 16] = np.inf
    xb = clf.to_device_colmajor(bad)
    st = clf.new_status()
    ob = clf.predict_proba_device(xb, n=1000, status=st)
    torch.cuda.synchronize()
    assert int(st[0].item()) & pkg.capi.ROW_NONFINITE
    assert torch.isnan(ob[77]).all() and torch.isnan(ob[500]).all()
    keep = np.setdiff1d(np.arange(1000), [77, 500])
    assert np.abs(ob.cpu().numpy()[keep] - so.predict_proba(arrays, X[:1000])[keep]).max() <= PROBA_TOL
    clf.close()


def test_tensor_core_engine_table_shapes(pkg, published, torch_cuda):
    """Support-vector counts around the block boundaries of the tensor-core kernel (1 to 4 blocks of up to 128 columns,
    the spare logistic column), the shared-memory limit (falls back to the SIMT loop), stacks without a logistic or a
    boosting member."""
    import copy
    rng = np.random.default_rng(17)
    base = published["arrays"]
    X = pkg.synth.host_rows(published["specs"], 61, 0, 9000).astype(np.float64)

    def with_svs(n_sv):
        arrays = copy.deepcopy(base)
        svc = arrays.members[0]
        pick = rng.integers(0, svc.sv.shape[0], n_sv)
        svc.sv = svc.sv[pick] + rng.normal(0, 0.05, (n_sv, 17))
        svc.dual_coef = svc.dual_coef[pick] * min(434.0 / n_sv, 4.0)
        order = np.argsort(svc.dual_coef > 0, kind="stable")
        svc.sv, svc.dual_coef = svc.sv[order], svc.dual_coef[order]
        n0 = int((svc.dual_coef <= 0).sum())
        svc.n_support = np.array([n0, n_sv - n0])
        return arrays

    for n_sv, engine in ((1, 2), (15, 2), (16, 2), (127, 2), (128, 2), (129, 2), (255, 2), (256, 2), (383, 2), (447, 2),
                         (448, 1), (600, 1)):
        arrays = with_svs(n_sv)
        clf = pkg.HFStackB200.from_arrays(arrays)
        assert clf.info.rbf_engine == engine, n_sv
        dev, lab = device_scores(clf, X)
        assert np.abs(dev - so.predict_proba(arrays, X)).max() <= PROBA_TOL, n_sv
        assert np.array_equal(arrays.classes[lab], so.predict(arrays, X)), n_sv
        clf.close()
    wm = pkg.synth_models.without_member
    for arrays in (wm(base, "linear"), wm(base, "gbt"), wm(wm(base, "gbt"), "linear")):
        clf = pkg.HFStackB200.from_arrays(arrays)
        assert clf.info.rbf_engine == 2
        dev, lab = device_scores(clf, X)
        assert np.abs(dev - so.predict_proba(arrays, X)).max() <= PROBA_TOL
        assert np.array_equal(arrays.classes[lab], so.predict(arrays, X))
        clf.close()


def test_streaming_pass_for_stacks_without_svc(pkg, published, retrained, torch_cuda):
    """Stacks without the SVC member go through hfe_stream_kernel (TMA-staged tiles, float32 links): same bars --
    probabilities <= 1e-6, labels and leaves bit-exact -- at ragged sizes, with a permuted column, NaN rows and the
    on-device accuracy count."""
    import torch
    wm = pkg.synth_models.without_member
    stacks = []
    for fx in (published, retrained):
        no_svc = wm(fx["arrays"], "svc")
        stacks += [(no_svc, fx["specs"]), (wm(no_svc, "gbt"), fx["specs"]), (wm(no_svc, "linear"), fx["specs"])]
    for arrays, specs in stacks:
        clf = pkg.HFStackB200.from_arrays(arrays)
        assert clf.info.n_support_vectors == 0
        X = pkg.synth.host_rows(specs, 41, 0, 40_000).astype(np.float64)
        want = so.predict_proba(arrays, X)
        got = clf.predict_proba(X)
        assert np.abs(got - want).max() <= PROBA_TOL
        assert np.array_equal(clf.predict(X), so.predict(arrays, X))
        assert np.abs(clf.decision_function(X) - so.decision_function(arrays, X)).max() <= 5e-6
        assert np.abs(clf.transform(X) - so.transform(arrays, X)).max() <= PROBA_TOL
        for n in (1, 3, 4, 5, 127, 128, 129, 511, 4099):
            assert np.abs(clf.predict_proba(X[:n]) - want[:n]).max() <= PROBA_TOL, n
            dev, dev_labels = device_scores(clf, X[:n])                      # the streaming kernel at this size
            assert np.abs(dev - want[:n]).max() <= PROBA_TOL, n
            assert np.array_equal(arrays.classes[dev_labels], so.predict(arrays, X[:n])), n
        bad = X[:300].copy()
        bad[77, 5] = np.nan
        with pytest.raises(ValueError):
            clf.predict_proba(bad)
        bad[77, 5] = np.inf
        with pytest.raises(ValueError):
            clf.predict_proba(bad)
        # one permutation-importance pass == scoring the materialised shuffle; accuracy counted on device
        n = 30_001
        perm = np.random.RandomState(3).permutation(n).astype(np.int32)
        col = 13
        Xp = X[:n].copy()
        Xp[:, col] = X[:n][perm, col]
        x, xp = clf.to_device_colmajor(X[:n]), clf.to_device_colmajor(Xp)
        y = so.predict(arrays, Xp).astype(np.uint8)
        y_dev = torch.from_numpy(y).cuda()
        out_a = torch.empty((n, 2), dtype=torch.float32, device="cuda")
        out_b = torch.empty_like(out_a)
        score = torch.zeros(1, dtype=torch.int64, device="cuda")
        clf.score_device(x, n=n, perm_col=col, perm_idx=torch.from_numpy(perm).cuda(), out=out_a, y=y_dev, score=score)
        clf.score_device(xp, n=n, out=out_b)
        torch.cuda.synchronize()
        assert torch.equal(out_a, out_b)
        assert int(score.item()) == n
        clf.close()


def test_integer_cohort_ingest(pkg, published, torch_cuda):
    """int64 (what predict_hf.py:29-36 passes) and uint8 cohort matrices through the device ingest kernels give the
    bytes the float64 route gives; the pinned pipeline accepts them too."""
    import torch
    clf = pkg.HFStackB200.from_arrays(published["arrays"])
    X = pkg.synth.host_rows(published["specs"], 31, 0, 4099)
    assert np.all(X == np.round(X)) and X.min() >= 0 and X.max() <= 255
    want = clf.to_device_colmajor(X.astype(np.float64))
    for dt in (np.int64, np.uint8, np.float32, np.int32):
        got = clf.to_device_colmajor(X.astype(dt))
        assert torch.equal(got, want), dt
    ref = clf.predict_proba(X.astype(np.float64))
    for tdt in (torch.int64, torch.uint8):
        xh = torch.from_numpy(X).to(tdt).pin_memory()
        out = clf.predict_proba_pinned(xh, chunk_rows=1024)
        assert np.abs(out.numpy().astype(np.float64) - ref).max() <= 1e-7
    big = np.array([[2 ** 40 + 1] * 17], dtype=np.int64)                  # int64 -> float64 -> float32, like numpy
    got = clf.to_device_colmajor(big)[:, 0].cpu().numpy()
    assert np.array_equal(got, big[0].astype(np.float64).astype(np.float32))
    with pytest.raises(ValueError):
        clf.predict_proba_pinned(torch.zeros((8, 17), dtype=torch.int16).pin_memory())
    clf.close()


def test_bounds_checked_build_on_small_and_ragged_cases(cuda_lib):
    """Every kernel at tiny / ragged sizes against the library built with -DHFE_BOUNDS_CHECK: any access outside the
    declared extents of x, perm_idx or the outputs traps and the run fails (stands in for compute-sanitizer)."""
    import os
    import subprocess
    import sys
    from conftest import ROOT
    checked = os.path.join(ROOT, "machine-learning-replications_b200", "libhfe_checked.so")
    assert os.path.exists(checked), "build() did not produce the bounds-checked library"
    env = dict(os.environ, HFE_LIB=checked)
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_target.py")], env=env,
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "sanitize target done" in res.stdout, res.stderr[-2000:]


def test_guard_probe_is_a_checked_contract(pkg, published, retrained, torch_cuda):
    """hfe_model_create scores a probe batch (support vectors in raw units and crossovers of them) through every fast
    engine and the float64 kernel; the handle reports what it saw and no band in use is below 8 x the observed error.
    A formula band made 1000 x too narrow on purpose is widened by the probe instead of letting labels flip, and a
    model whose dual coefficients are 1000 x larger (errors 1000 x larger) keeps its labels too."""
    import copy
    torch = torch_cuda
    for case in (published, retrained):
        clf = pkg.HFStackB200(case["blob"], device=0)
        info = clf.info
        assert info.probe_rows == 65536 and info.guard_widened == 0
        assert 0.0 < info.svc_observed_err and 8.0 * info.svc_observed_err <= info.svc_guard
        assert 0.0 < info.label_observed_err and 8.0 * info.label_observed_err <= info.label_guard
        print(f"{case['name']}: svc_guard {info.svc_guard:.3e} / observed {info.svc_observed_err:.3e} = "
              f"{info.svc_guard / info.svc_observed_err:.0f}x; label_guard {info.label_guard:.3e} / observed "
              f"{info.label_observed_err:.3e} = {info.label_guard / info.label_observed_err:.0f}x")
        wide = info.svc_guard
        # the formula made far too optimistic: the probe has to put the band back above what it measures
        clf.calibrate_guards(1e-3)
        info = clf.info
        assert info.guard_widened & 1 and info.svc_guard == 8.0 * info.svc_observed_err and info.svc_guard < wide
        assert 8.0 * info.label_observed_err <= info.label_guard
        n = 2_000_000
        x = pkg.synth_rows_device(case["specs"], 41, 0, n)
        lab = torch.empty(n, dtype=torch.uint8, device="cuda")
        lab64 = torch.empty(n, dtype=torch.uint8, device="cuda")
        for engine in ("simt", "tensor"):
            clf.set_rbf_engine(engine)
            clf.score_device(x, n=n, labels=lab)
            clf.score_f64_device(x, n=n, labels=lab64)
            torch.cuda.synchronize()
            assert torch.equal(lab, lab64), engine
        clf.set_rbf_engine("auto")
        clf.calibrate_guards(1.0)
        assert clf.info.guard_widened == 0 and clf.info.svc_guard == wide
        clf.close()
    # adversarial table: dual coefficients (and intercept) x 1000 -> decision values and their float32 error x 1000
    arrays = copy.deepcopy(published["arrays"])
    svc = arrays.members[0]
    svc.dual_coef = svc.dual_coef * 1e3
    svc.intercept = svc.intercept * 1e3
    svc.prob_a = svc.prob_a / 1e3                         # keeps the Platt argument, hence the score distribution
    big = pkg.HFStackB200.from_arrays(arrays)
    ref = pkg.HFStackB200(published["blob"], device=0)
    assert big.info.svc_observed_err > 100.0 * ref.info.svc_observed_err
    assert 8.0 * big.info.svc_observed_err <= big.info.svc_guard
    X = pkg.synth.host_rows(published["specs"], 43, 0, 20000).astype(np.float64)
    assert np.array_equal(big.predict(X), so.predict(arrays, X))
    dev, dev_labels = device_scores(big, X)
    assert np.abs(dev - so.predict_proba(arrays, X)).max() <= PROBA_TOL
    assert np.array_equal(big.classes_[dev_labels], so.predict(arrays, X))
    big.close()
    ref.close()
