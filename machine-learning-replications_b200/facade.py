"""Host-side mirror of the reference's model object: the scikit-learn estimator protocol over the CUDA scorer.

``predict_hf.py:36`` and ``train_ensemble_public.py:62`` call ``clf.predict_proba(X)`` on an unpickled
``StackingClassifier``.  ``HFStackB200`` is the object those lines get instead: same method names, argument meaning,
return types (float64 ``[n,2]`` ordered as ``classes_``; ``predict`` -> float64 labels decided by the combiner's strict
``z > 0``, SK/linear_model/_base.py:398-425) and the same ``ValueError`` behaviour for malformed input
(1-D input, wrong width, empty, NaN/Inf -- SK/svm/_base.py:648-660).  Everything numeric happens in ``libhfe.so``;
there is no CPU path.

numpy entry points go through ``hfe_predict_host`` (H2D copy -> device ingest/transpose -> fused kernel -> D2H).  The
``*_device`` methods take a column-major float32 CUDA tensor ``[17, ld]`` (torch is used for device memory and
streams only) and enqueue on the current stream without synchronising.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .export import ModelArrays, arrays_from_published_pickle, arrays_from_stack, read_blob, write_blob
from .schema import N_FEATURES


def _validate_X(X, n_features, check_finite=True):
    """The checks scikit-learn's ``validate_data`` applies on this path, with its messages."""
    arr = np.asarray(X)
    if arr.dtype == object or arr.dtype.kind in "USV":
        arr = arr.astype(np.float64)                      # raises ValueError for non-numeric content, like sklearn
    if arr.dtype.kind == "c":
        raise ValueError(f"Complex data not supported\n{arr}\n")
    if arr.ndim == 0:
        raise ValueError(f"Expected 2D array, got scalar array instead:\narray={arr!r}.")
    if arr.ndim == 1:
        raise ValueError(f"Expected 2D array, got 1D array instead:\narray={arr!r}.\nReshape your data either using "
                         "array.reshape(-1, 1) if your data has a single feature or array.reshape(1, -1) if it "
                         "contains a single sample.")
    if arr.ndim > 2:
        raise ValueError(f"Found array with dim {arr.ndim}, while dim <= 2 is required.")
    if arr.shape[0] == 0:
        raise ValueError(f"Found array with 0 sample(s) (shape={arr.shape}) while a minimum of 1 is required.")
    if arr.shape[1] != n_features:
        raise ValueError(f"X has {arr.shape[1]} features, but HFStackB200 is expecting {n_features} features as "
                         "input.")
    arr = np.ascontiguousarray(arr, dtype=np.float64)
    if check_finite:
        _raise_if_not_finite(arr)
    return arr


class _ResultPool:
    """Large result arrays are carved out of recycled buffers.  A fresh 160 MB array costs the caller's threads 40 000
    page faults (and the kernel as many zeroed pages) every call -- on the 16-CPU box that was 12.8 of the 27.9 ms a
    10 M-row ``predict_proba`` took.  Here the buffer of a result goes back to a small pool when the array handed out
    (and every view of it) has been garbage-collected, and the next call of the same size writes into pages that are
    already mapped.  A buffer that any array or view of the caller's still refers to is never reused; nothing is shared
    between live results."""

    MIN_BYTES = 1 << 22            # below this, numpy's own allocator is as good
    MAX_HELD = 2                   # buffers kept waiting for reuse
    MAX_HELD_BYTES = 1 << 30

    def __init__(self):
        import threading
        self._free = []            # [(nbytes, owner uint8 array)]
        self._lock = threading.Lock()

    def _give_back(self, owner):
        import sys
        with self._lock:
            if len(self._free) >= self.MAX_HELD:
                # full: let go of the oldest buffer a caller still holds on to (it stays theirs; it just will not be
                # recycled), so results that live long do not clog the pool
                for k in range(len(self._free)):
                    if sys.getrefcount(self._free[k][1]) > 2:
                        self._free.pop(k)
                        break
            held = sum(b for b, _ in self._free)
            if len(self._free) < self.MAX_HELD and held + owner.nbytes <= self.MAX_HELD_BYTES:
                self._free.append((owner.nbytes, owner))

    def empty(self, shape, dtype):
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        if nbytes < self.MIN_BYTES:
            return np.empty(shape, dtype=dtype)
        import sys
        owner = None
        with self._lock:
            for k in range(len(self._free)):
                # a view the caller still holds has the buffer itself as its base (numpy collapses base chains): such a
                # buffer stays where it is until the tuple in this list is its only reference
                if self._free[k][0] == nbytes and sys.getrefcount(self._free[k][1]) == 2:
                    owner = self._free.pop(k)[1]
                    break
        if owner is None:
            owner = np.empty(nbytes, dtype=np.uint8)
        # numpy collapses base chains: the array handed out and every view of it keep `owner` alive, so its reference count
        # tells whether the caller still uses the buffer; it is listed for reuse right away and taken when that count drops
        out = owner.view(dtype).reshape(shape)
        self._give_back(owner)
        return out


_results = _ResultPool()


def _raise_if_not_finite(arr):
    """scikit-learn's finiteness check with its messages.  The scoring entry points skip this host-side scan (it costs
    as much as the whole device call on large arrays): the device pass flags non-finite values, and only then is the
    array scanned here to word the error like scikit-learn does."""
    if not np.isfinite(arr).all():
        if np.isnan(arr).any():
            raise ValueError("Input X contains NaN.")
        raise ValueError("Input X contains infinity or a value too large for dtype('float64').")


class HFStackB200:
    """Drop-in for the fitted ``StackingClassifier`` of train_ensemble_public.py:43-48 on one B200."""

    def __init__(self, blob: bytes, device: int = 0, arrays: ModelArrays | None = None):
        self._lib = capi.load_library()
        self._blob = bytes(blob)
        self._arrays = arrays if arrays is not None else read_blob(self._blob)
        self.device = int(device)
        handle = C.c_void_p()
        capi.check(self._lib.hfe_model_create(self._blob, len(self._blob), self.device, C.byref(handle)))
        self._handle = handle
        self.classes_ = np.array(self._arrays.classes, dtype=np.float64)
        self.n_features_in_ = N_FEATURES
        info = capi.ModelInfo()
        capi.check(self._lib.hfe_model_get_info(self._handle, C.byref(info)))
        self.info = info
        self.n_members = int(info.n_members)
        self.last_status = None

    # -- constructors ----------------------------------------------------------------------------------------------
    @classmethod
    def from_arrays(cls, arrays: ModelArrays, device: int = 0):
        return cls(write_blob(arrays), device=device, arrays=arrays)

    @classmethod
    def from_sklearn(cls, stack, device: int = 0):
        """Walk a live fitted StackingClassifier (attributes only) -- replaces pickling it."""
        return cls.from_arrays(arrays_from_stack(stack, source="live sklearn object"), device=device)

    @classmethod
    def from_published_pickle(cls, path, device: int = 0):
        """Statically decode the reference's hf_predict_model.pkl (never unpickled) -- replaces predict_hf.py:33-34."""
        return cls.from_arrays(arrays_from_published_pickle(path), device=device)

    @classmethod
    def from_blob_file(cls, path, device: int = 0):
        with open(path, "rb") as fh:
            return cls(fh.read(), device=device)

    def set_rbf_engine(self, engine):
        """``"auto"`` / ``"simt"`` / ``"tensor"``: what evaluates the RBF sum of the SVC member (both meet the same bars)."""
        code = {"auto": 0, "simt": 1, "tensor": 2}[engine]
        capi.check(self._lib.hfe_model_set_rbf_engine(self._handle, code))
        capi.check(self._lib.hfe_model_get_info(self._handle, C.byref(self.info)))
        return self

    def calibrate_guards(self, formula_scale: float = 1.0):
        """Re-run the creation-time guard probe (``hfe_model_calibrate_guards``) with the formula bands scaled by
        ``formula_scale``; ``self.info`` then holds the bands in use and the errors the probe observed."""
        capi.check(self._lib.hfe_model_calibrate_guards(self._handle, float(formula_scale)))
        capi.check(self._lib.hfe_model_get_info(self._handle, C.byref(self.info)))
        return self

    def set_tree_engine(self, engine):
        """``"auto"`` / ``"l1"`` / ``"smem"`` / ``"l1+l2window"``: how general boosted trees are walked (same results)."""
        capi.check(self._lib.hfe_model_set_tree_engine(self._handle, {"auto": 0, "l1": 1, "smem": 2, "l1+l2window": 3}[engine]))
        return self

    def close(self):
        if getattr(self, "_handle", None) is not None and self._handle:
            self._lib.hfe_model_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- numpy (host) protocol -------------------------------------------------------------------------------------
    def _host_call(self, X, want_labels=False, want_decision=False, want_meta=False):
        arr = _validate_X(X, self.n_features_in_, check_finite=False)
        n = arr.shape[0]
        proba = _results.empty((n, 2), np.float64)
        labels = np.empty(n, dtype=np.uint8) if want_labels else None
        decision = np.empty(n, dtype=np.float64) if want_decision else None
        meta = np.empty((n, self.n_members), dtype=np.float64) if want_meta else None

        def ptr(a):
            return a.ctypes.data_as(C.c_void_p) if a is not None else None

        code = self._lib.hfe_predict_host(self._handle, ptr(arr), n, ptr(proba), ptr(labels), ptr(decision),
                                          ptr(meta))
        if code == capi.HFE_ERR_NONFINITE_INPUT:
            _raise_if_not_finite(arr)
            raise ValueError("Input X contains infinity or a value too large for dtype('float32').")
        capi.check(code)
        return proba, labels, decision, meta

    def predict_proba(self, X):
        """``[n,2]`` float64, columns ordered as ``classes_`` (predict_hf.py:36; train_ensemble_public.py:62)."""
        return self._host_call(X)[0]

    def predict(self, X):
        """``classes_[z > 0]`` (SK/ensemble/_stacking.py:711-766)."""
        return self.classes_[self._host_call(X, want_labels=True)[1].astype(np.int64)]

    def decision_function(self, X):
        return self._host_call(X, want_decision=True)[2]

    def transform(self, X):
        """Member P(class 1) columns ``[n, n_members]`` (SK/ensemble/_stacking.py:814-830)."""
        return self._host_call(X, want_meta=True)[3]

    def predict_log_proba(self, X):
        return np.log(self.predict_proba(X))

    def score(self, X, y):
        """Mean accuracy, the default scorer permutation importance uses (ClassifierMixin.score)."""
        return float(np.mean(self.predict(X) == np.asarray(y)))

    def apply(self, X):
        """Leaf ids of the boosting member ``[n, n_trees]`` (GradientBoostingClassifier.apply, ids as integers)."""
        import torch

        arr = _validate_X(X, self.n_features_in_)
        x = self.to_device_colmajor(arr)
        n = arr.shape[0]
        leaves = torch.empty((n, int(self.info.n_trees)), dtype=torch.int16, device=x.device)
        capi.check(self._lib.hfe_apply(self._handle, x.data_ptr(), n, x.shape[1], leaves.data_ptr(), self._stream()))
        return leaves.cpu().numpy().view(np.uint16).astype(np.int64)

    # -- device-native protocol (timed path) ---------------------------------------------------------------------------
    def _stream(self):
        import torch
        return torch.cuda.current_stream(self.device).cuda_stream

    def to_device_colmajor(self, X):
        """Row-major host array -> column-major float32 CUDA tensor ``[17, ld]`` via the device ingest kernel."""
        import torch

        arr = np.asarray(X)
        if arr.dtype not in (np.float64, np.float32, np.int64, np.uint8):
            arr = arr.astype(np.float64)
        arr = np.ascontiguousarray(arr)
        n = arr.shape[0]
        ld = (n + 3) & ~3
        with torch.cuda.device(self.device):
            src = torch.from_numpy(arr).to(f"cuda:{self.device}")
            dst = torch.zeros((self.n_features_in_, ld), dtype=torch.float32, device=src.device)
            status = torch.zeros(capi.STATUS_HEADER_WORDS, dtype=torch.int32, device=src.device)
            capi.check(self._ingest_for(src.dtype)(src.data_ptr(), n, self.n_features_in_, dst.data_ptr(), ld,
                                                   status.data_ptr(), self._stream()))
        return dst

    def _ingest_for(self, dtype):
        """The device ingest entry point for a row-major matrix of torch dtype ``dtype``."""
        import torch
        table = {torch.float64: self._lib.hfe_ingest_rowmajor_f64, torch.float32: self._lib.hfe_ingest_rowmajor_f32,
                 torch.int64: self._lib.hfe_ingest_rowmajor_i64, torch.uint8: self._lib.hfe_ingest_rowmajor_u8}
        if dtype not in table:
            raise ValueError("row-major cohort matrices must be float64, float32, int64 or uint8")
        return table[dtype]

    def new_status(self):
        """Per-call status + work-list buffer (256 KiB of device memory; calls in flight at once need one each)."""
        import torch
        return torch.zeros(capi.STATUS_WORDS, dtype=torch.int32, device=f"cuda:{self.device}")

    def predict_proba_device(self, x, n=None, out=None, status=None):
        """Enqueue the fused kernel on the current stream.  ``x``: float32 CUDA tensor ``[17, ld]`` (column-major
        feature matrix); returns the float32 ``[n,2]`` output tensor.  Does not synchronise."""
        import torch

        n = x.shape[1] if n is None else int(n)
        if out is None:
            out = torch.empty((n, 2), dtype=torch.float32, device=x.device)
        if status is None:
            status = self.new_status()
        self._lazy_check()
        capi.check(self._lib.hfe_predict_proba(self._handle, x.data_ptr(), n, x.stride(0), out.data_ptr(),
                                               status.data_ptr(), self._stream()))
        self.last_status = status
        # an event behind the call: check() / the next device call look at the call's error bits once it has passed
        # (at the status array the caller gave -- a caller who reuses it for another call sees that call's bits)
        if getattr(self, "_pending", None) is None:
            self._pending = torch.cuda.Event()
        self._pending.record(torch.cuda.current_stream(self.device))
        self._pending_live = True
        return out

    def _raise_for_flags(self, flags):
        if flags & capi.ROW_NONFINITE:
            raise ValueError("Input X contains NaN or infinity.")
        if flags & capi.WORKLIST_OVERFLOW:
            raise capi.HfeError(capi.HFE_ERR_TOO_LARGE,
                                "work list overflow: more than WORKLIST_CAPACITY rows of the last device call needed "
                                "the float64 re-evaluation and the rows beyond it kept their float32 result; rerun "
                                "that batch through predict_proba_f64_device")

    def _lazy_check(self):
        """If the previous ``predict_proba_device`` call has finished, look at its error bits (no waiting).  A work-list
        overflow of that call is reported as a ``RuntimeWarning`` here -- the caller never asked about it, and its
        non-finite rows are visible as NaN outputs anyway; ``check()`` raises."""
        if getattr(self, "_pending_live", False) and self._pending.query():
            self._pending_live = False
            if int(self.last_status[0].item()) & capi.WORKLIST_OVERFLOW:
                import warnings
                warnings.warn("the previous predict_proba_device call overflowed its work list: rows beyond "
                              "WORKLIST_CAPACITY kept their float32-path result; rerun that batch through "
                              "predict_proba_f64_device (clf.check() raises for this)", RuntimeWarning, stacklevel=3)

    def check(self, status=None):
        """Wait for the device calls issued so far and raise for their error bits: ``ValueError`` for NaN/Inf input
        (what the numpy entry points raise), ``HfeError`` for a work-list overflow.  ``status``: a status tensor of a
        finished call to look at instead of the last ``predict_proba_device`` call's snapshot."""
        import torch
        torch.cuda.current_stream(self.device).synchronize()
        if status is not None:
            self._raise_for_flags(int(status[0].item()))
        elif getattr(self, "last_status", None) is not None:
            self._pending_live = False
            self._raise_for_flags(int(self.last_status[0].item()))

    def score_device(self, x, n=None, perm_col=-1, perm_idx=None, out=None, labels=None, decision=None, meta=None,
                     y=None, score=None, status=None):
        """``hfe_score``: any subset of outputs; optional permuted column and on-device accuracy count."""
        n = x.shape[1] if n is None else int(n)
        if status is None:
            status = self.new_status()

        def p(t):
            return t.data_ptr() if t is not None else None

        capi.check(self._lib.hfe_score(self._handle, x.data_ptr(), n, x.stride(0), int(perm_col), p(perm_idx), p(out),
                                       p(labels), p(decision), p(meta), p(y), p(score), status.data_ptr(),
                                       self._stream()))
        self.last_status = status
        return status

    def predict_proba_pinned(self, x_host, out_host=None, chunk_rows=1 << 21):
        """End-to-end host path for large cohorts: ``x_host`` is a row-major float32, float64, int64 or uint8 ``[n, 17]``
        torch tensor in pinned host memory (uint8: every schema variable is a small non-negative integer, and 17 bytes
        per patient instead of 68 cross PCIe); returns a pinned float32 ``[n, 2]`` tensor.  Rows are cut into chunks that
        alternate between two streams, so the H2D copy of one chunk overlaps the ingest + fused kernel + D2H copy of
        the other.  Raises ``ValueError`` for NaN/Inf like the numpy entry points.  Synchronises before returning."""
        import torch

        n = int(x_host.shape[0])
        if x_host.dim() != 2 or x_host.shape[1] != self.n_features_in_:
            raise ValueError(f"X has {tuple(x_host.shape)} shape, expected [n, {self.n_features_in_}]")
        if not x_host.is_contiguous():
            raise ValueError("x_host must be a contiguous tensor")
        ingest = self._ingest_for(x_host.dtype)
        if out_host is None:
            out_host = torch.empty((n, 2), dtype=torch.float32).pin_memory()
        dev = f"cuda:{self.device}"
        chunk = int(min(max(chunk_rows, 4), max(n, 4)) + 3) & ~3
        key = (chunk, x_host.dtype)
        if getattr(self, "_pipe_key", None) != key:
            with torch.cuda.device(self.device):
                self._pipe = [dict(stream=torch.cuda.Stream(device=dev),
                                   raw=torch.empty((chunk, self.n_features_in_), dtype=x_host.dtype, device=dev),
                                   col=torch.empty((self.n_features_in_, chunk), dtype=torch.float32, device=dev),
                                   out=torch.empty((chunk, 2), dtype=torch.float32, device=dev),
                                   st_in=torch.zeros(capi.STATUS_HEADER_WORDS, dtype=torch.int32, device=dev),
                                   st=torch.zeros(capi.STATUS_WORDS, dtype=torch.int32, device=dev),
                                   flags=torch.zeros(1, dtype=torch.int32, device=dev)) for _ in range(2)]
            self._pipe_key = key
        for slot in self._pipe:
            slot["flags"].zero_()
        torch.cuda.current_stream(self.device).synchronize()
        for i, begin in enumerate(range(0, n, chunk)):
            rows = min(chunk, n - begin)
            slot = self._pipe[i & 1]
            s = slot["stream"]
            with torch.cuda.stream(s):
                slot["raw"][:rows].copy_(x_host[begin:begin + rows], non_blocking=True)
                capi.check(ingest(slot["raw"].data_ptr(), rows, self.n_features_in_, slot["col"].data_ptr(), chunk,
                                  slot["st_in"].data_ptr(), s.cuda_stream))
                capi.check(self._lib.hfe_predict_proba(self._handle, slot["col"].data_ptr(), rows, chunk,
                                                       slot["out"].data_ptr(), slot["st"].data_ptr(), s.cuda_stream))
                slot["flags"] |= slot["st_in"][:1] | slot["st"][:1]
                out_host[begin:begin + rows].copy_(slot["out"][:rows], non_blocking=True)
        bad = 0
        for slot in self._pipe:
            slot["stream"].synchronize()
            bad |= int(slot["flags"].item())
        if bad & capi.ROW_NONFINITE:
            raise ValueError("Input X contains NaN or infinity.")
        if bad & capi.WORKLIST_OVERFLOW:
            # more rows than one call's work list holds needed float64: redo the batch on the all-float64 path
            slot = self._pipe[0]
            for begin in range(0, n, chunk):
                rows = min(chunk, n - begin)
                slot["raw"][:rows].copy_(x_host[begin:begin + rows])
                capi.check(ingest(slot["raw"].data_ptr(), rows, self.n_features_in_, slot["col"].data_ptr(), chunk,
                                  slot["st_in"].data_ptr(), self._stream()))
                o64 = self.predict_proba_f64_device(slot["col"], n=rows)
                out_host[begin:begin + rows].copy_(o64.float())
            torch.cuda.current_stream(self.device).synchronize()
        return out_host

    def predict_proba_f64_device(self, x, n=None, out=None, status=None):
        """Every row through the float64 path (calibration aid)."""
        import torch

        n = x.shape[1] if n is None else int(n)
        if out is None:
            out = torch.empty((n, 2), dtype=torch.float64, device=x.device)
        if status is None:
            status = self.new_status()
        capi.check(self._lib.hfe_predict_proba_f64(self._handle, x.data_ptr(), n, x.stride(0), out.data_ptr(),
                                                   status.data_ptr(), self._stream()))
        return out

    def score_f64_device(self, x, n=None, out=None, labels=None, decision=None, meta=None, status=None):
        """``hfe_score_f64``: any subset of float64-path outputs (the parity harness's exhaustive label check)."""
        n = x.shape[1] if n is None else int(n)
        if status is None:
            status = self.new_status()

        def p(t):
            return t.data_ptr() if t is not None else None

        capi.check(self._lib.hfe_score_f64(self._handle, x.data_ptr(), n, x.stride(0), p(out), p(labels), p(decision),
                                           p(meta), status.data_ptr(), self._stream()))
        return status

    def apply_device(self, x, n=None):
        import torch

        n = x.shape[1] if n is None else int(n)
        leaves = torch.empty((n, int(self.info.n_trees)), dtype=torch.int16, device=x.device)
        capi.check(self._lib.hfe_apply(self._handle, x.data_ptr(), n, x.stride(0), leaves.data_ptr(), self._stream()))
        return leaves


def synth_rows_device(specs, seed, row0, n, device=0, out=None):
    """Column-major float32 ``[len(specs), ld]`` synthetic rows generated on the device (``hfe_synth_rows``)."""
    import torch

    from .synth import pack_specs

    lib = capi.load_library()
    col_id, base, step, off, thr = pack_specs(specs)
    dev = f"cuda:{device}"
    ld = (n + 3) & ~3
    with torch.cuda.device(device):
        t_col = torch.from_numpy(col_id).to(dev)
        t_base = torch.from_numpy(base).to(dev)
        t_step = torch.from_numpy(step).to(dev)
        t_off = torch.from_numpy(off).to(dev)
        t_thr = torch.from_numpy(thr.view(np.int32)).to(dev)
        if out is None:
            out = torch.zeros((len(specs), ld), dtype=torch.float32, device=dev)
        capi.check(lib.hfe_synth_rows(int(seed), int(row0), int(n), len(specs), t_col.data_ptr(), t_base.data_ptr(),
                                      t_step.data_ptr(), t_off.data_ptr(), t_thr.data_ptr(), out.data_ptr(),
                                      out.stride(0), torch.cuda.current_stream(device).cuda_stream))
        torch.cuda.current_stream(device).synchronize()     # the table tensors above must outlive the kernel
    return out
