"""Host-side pieces of permutation importance and the evaluation metrics (no GPU)."""
import numpy as np


def test_shuffled_indices_compose_like_scikit_learn(pkg):
    """SK/inspection/_permutation_importance.py:66-74 keeps one index vector, shuffles it again for every repeat and
    applies it to the *already permuted* column; the effective source index of repeat r must reproduce that."""
    n, repeats, seed = 1000, 4, 12345
    eff = pkg.inspection.shuffled_indices(n, repeats, seed)
    col = np.random.default_rng(0).normal(size=n)
    rng = np.random.RandomState(seed)
    idx = np.arange(n)
    permuted = col.copy()
    for r in range(repeats):
        rng.shuffle(idx)
        permuted = permuted[idx]
        assert np.array_equal(col[eff[r]], permuted)
        assert sorted(eff[r].tolist()) == list(range(n))


def test_binary_report_and_curves_match_scikit_learn(pkg):
    from sklearn import metrics as skm
    rng = np.random.default_rng(1)
    n = 5000
    y = (rng.random(n) < 0.3).astype(int)
    bins = pkg.capi.METRIC_BINS if hasattr(pkg.capi, "METRIC_BINS") else 65536
    p = np.clip(0.3 * y + 0.7 * rng.random(n), 0, 1).astype(np.float32)
    q = np.floor(p.astype(np.float64) * (bins - 1) + 0.5).astype(np.int64)           # the kernel's binning
    hist = np.zeros((2, bins), dtype=np.int64)
    np.add.at(hist, (y, q), 1)
    pred = p > 0.5
    conf = np.array([[np.sum(~pred & (y == 0)), np.sum(pred & (y == 0))],
                     [np.sum(~pred & (y == 1)), np.sum(pred & (y == 1))]])
    rep = pkg.metrics.binary_report(conf)
    ref = skm.classification_report(y, pred.astype(int), output_dict=True)
    for cls in ("0", "1"):
        for key in ("precision", "recall", "f1-score", "support"):
            assert abs(rep[cls][key] - ref[cls][key]) < 1e-12
    assert abs(rep["accuracy"] - ref["accuracy"]) < 1e-12
    binned = q / (bins - 1)
    assert abs(pkg.metrics.roc_auc(hist) - skm.roc_auc_score(y, binned)) < 1e-9
    assert abs(pkg.metrics.average_precision(hist) - skm.average_precision_score(y, binned)) < 1e-9


def test_native_shuffles_are_numpys(pkg, cuda_lib):
    """hfe_perm_indices restates RandomState(seed).shuffle (legacy MT19937, masked rejection) and scikit-learn's cumulative
    use of it (SK/inspection/_permutation_importance.py:66-74): bit for bit the numpy restatement, hence scikit-learn's."""
    insp = pkg.inspection
    for n, reps, seed in ((0, 2, 1), (1, 3, 5), (2, 2, 7), (3, 4, 0), (255, 3, 9), (256, 3, 9), (257, 5, 4294967295), (513, 2, 77), (50_000, 5, 2020), (65_537, 2, 123456789)):
        assert np.array_equal(insp.native_shuffled_indices(n, reps, seed), insp.shuffled_indices(n, reps, seed)), (n, seed)


def test_permutation_passes_refuses_bad_job_lists(pkg, cuda_lib):
    """hfe_permutation_passes validates its job list before it touches a device: repeats must ascend and stay in range."""
    import ctypes as C
    lib = cuda_lib
    cols = (C.c_int32 * 3)(0, 1, 2)
    host = np.zeros((2, 8), dtype=np.int32)
    dummy = C.c_void_p(16)                                    # never dereferenced: validation comes first

    def call(reps, n_repeats=2, n_jobs=3):
        r = (C.c_int32 * 3)(*reps)
        return lib.hfe_permutation_passes(None, dummy, 8, 8, dummy, 1, n_repeats, n_jobs, cols, r,
                                          host.ctypes.data_as(C.c_void_p), dummy, dummy, None, dummy, None)

    assert call((1, 0, 1)) == pkg.capi.HFE_ERR_INVALID_ARGUMENT          # descending repeat
    assert call((0, 1, 2)) == pkg.capi.HFE_ERR_INVALID_ARGUMENT          # repeat out of range
    assert call((0, 0, 1), n_jobs=-1) == pkg.capi.HFE_ERR_INVALID_ARGUMENT
    assert b"ascending" in lib.hfe_last_error_string() or b"job list" in lib.hfe_last_error_string()
