"""The allocation-free entry point (collected last on purpose: a run that stops at the first failure has then already
run every other test).

``hfe_predict_proba_noalloc`` exists for callers that are handed their input and output buffers by a host process
that owns all device memory (a scoring plug-in inside a database or inference server, a service that may not allocate
once it is serving): they cannot provide the 256 KiB status / work-list array the ordinary entry point wants, so the
library keeps that array in its own image.  Here it is called through the C ABI twice in a row on different batches
and compared with the ordinary entry point on the same rows; the second call must not be affected by what the first
one left in the shared block, and the call must not change the amount of free device memory.
"""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
PROBA_TOL = 1e-6          # the tolerance the parity tests use (float32 probabilities)


def test_noalloc_entry_point_matches_ordinary_entry_point(pkg, cuda_lib, published):
    import torch
    assert torch.cuda.is_available()
    lib = cuda_lib
    clf = pkg.HFStackB200(published["blob"], device=0)
    # a batch that is known to put rows on the work list (the solver-edge / label-edge rows of the fixture), padded
    # with ordinary synthetic rows, and a second, different batch of another size
    g = published["golden"]
    first = np.concatenate([g["X"], pkg.synth.host_rows(published["specs"], 31, 0, 200_000)]).astype(np.float64)
    second = pkg.synth.host_rows(published["specs"], 32, 5_000_000, 65_537).astype(np.float64)
    for batch in (first, second):
        n = batch.shape[0]
        x = clf.to_device_colmajor(batch)
        want = torch.empty((n, 2), dtype=torch.float32, device="cuda")
        status = clf.new_status()
        pkg.capi.check(lib.hfe_predict_proba(clf._handle, x.data_ptr(), n, x.stride(0), want.data_ptr(),
                                             status.data_ptr(), None))
        got = torch.full((n, 2), -1.0, dtype=torch.float32, device="cuda")
        flags = torch.full((1,), 77, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        pkg.capi.check(lib.hfe_predict_proba_noalloc(clf._handle, x.data_ptr(), n, x.stride(0), got.data_ptr(),
                                                     flags.data_ptr(), None))
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before            # the call took no device memory
        assert int(flags.item()) == int(status[0].item()) == 0
        assert float((got - want).abs().max().item()) <= PROBA_TOL
        assert torch.equal(got, want)                                  # same kernels, same rows: identical bits
    assert int(status[1].item()) >= 0
    # the fixture's edge rows really did go through the work list on the first batch
    x = clf.to_device_colmajor(first)
    st = clf.new_status()
    clf.predict_proba_device(x, n=first.shape[0], status=st)
    torch.cuda.synchronize()
    assert int(st[1].item()) > 0


def test_noalloc_entry_point_limits_and_errors(pkg, cuda_lib, published):
    import torch
    lib = cuda_lib
    clf = pkg.HFStackB200(published["blob"], device=0)
    x = torch.zeros((17, 64), dtype=torch.float32, device="cuda")
    x[6] = 1.0
    out = torch.empty((64, 2), dtype=torch.float32, device="cuda")
    too_many = pkg.capi.NOALLOC_MAX_ROWS + 1
    rc = lib.hfe_predict_proba_noalloc(clf._handle, x.data_ptr(), too_many, too_many + 3 & ~3, out.data_ptr(), None,
                                       None)
    assert rc == 6 and b"HFE_NOALLOC_MAX_ROWS" in lib.hfe_last_error_string()       # HFE_ERR_TOO_LARGE
    assert lib.hfe_predict_proba_noalloc(None, x.data_ptr(), 64, 64, out.data_ptr(), None, None) == 1
    # non-finite input is reported through the caller's flags word, and flags may be omitted
    x[3, 10] = float("nan")
    flags = torch.zeros(1, dtype=torch.int32, device="cuda")
    pkg.capi.check(lib.hfe_predict_proba_noalloc(clf._handle, x.data_ptr(), 64, 64, out.data_ptr(), flags.data_ptr(),
                                                 None))
    torch.cuda.synchronize()
    assert int(flags.item()) & pkg.capi.ROW_NONFINITE
    assert torch.isnan(out[10]).all() and torch.isfinite(out[11]).all()
    x[3, 10] = 0.0
    pkg.capi.check(lib.hfe_predict_proba_noalloc(clf._handle, x.data_ptr(), 64, 64, out.data_ptr(), None, None))
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
