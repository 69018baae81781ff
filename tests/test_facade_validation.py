"""Input validation of the estimator facade mirrors scikit-learn's errors.  CPU only (no model handle is created)."""
import numpy as np
import pytest


def test_validate_X_errors(pkg):
    v = pkg.facade._validate_X
    with pytest.raises(ValueError, match="Expected 2D array, got 1D"):
        v(np.zeros(17), 17)
    with pytest.raises(ValueError, match="scalar"):
        v(3.0, 17)
    with pytest.raises(ValueError, match="0 sample"):
        v(np.zeros((0, 17)), 17)
    with pytest.raises(ValueError, match="16 features"):
        v(np.zeros((4, 16)), 17)
    bad = np.zeros((3, 17))
    bad[1, 2] = np.nan
    with pytest.raises(ValueError, match="NaN"):
        v(bad, 17)
    bad[1, 2] = np.inf
    with pytest.raises(ValueError, match="infinity"):
        v(bad, 17)
    with pytest.raises(ValueError):
        v(np.array([["a"] * 17]), 17)


def test_validate_X_accepts_what_the_reference_passes(pkg):
    v = pkg.facade._validate_X
    ints = np.reshape([1, 1, 0, 0, 1, 0, 1, 1, 0, 0, 0, 0, 0, 13, 0, 0, 55], (1, -1))       # predict_hf.py:36
    out = v(ints, 17)
    assert out.dtype == np.float64 and out.flags.c_contiguous and out.shape == (1, 17)
    f = np.asfortranarray(np.random.default_rng(0).random((5, 17)).astype(np.float32))
    assert np.array_equal(v(f, 17), f.astype(np.float64))
    assert v([[0.0] * 17], 17).shape == (1, 17)


def test_missing_library_fails_loudly(pkg, tmp_path, monkeypatch):
    monkeypatch.setattr(pkg.capi, "_lib", None)
    with pytest.raises(FileNotFoundError, match="no CPU fallback"):
        pkg.capi.load_library(str(tmp_path / "libhfe.so"))
