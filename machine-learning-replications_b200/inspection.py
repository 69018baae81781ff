"""Permutation feature importance on the device (BASELINE.json config 5).

Same procedure as ``sklearn.inspection.permutation_importance`` (SK/inspection/_permutation_importance.py:31-76,
278-313) with its default scorer for classifiers (accuracy): one ``random_seed`` is drawn from ``random_state`` and
*every* column restarts a ``RandomState(random_seed)``, shuffles an index vector cumulatively ``n_repeats`` times,
applies it each time to the already permuted column, and scores the matrix.  Here the ``n_repeats`` index vectors are made
once on the host (so they are exactly scikit-learn's), uploaded as int32, and each (column, repeat) pass is one launch
of the fused kernel with ``perm_col`` / ``perm_idx`` -- the shuffled matrix is never materialised -- and an on-device
count of ``label == y``.  With several ranks the 17 x n_repeats passes are dealt in contiguous blocks and the counts are summed
with one all-reduce of 86 integers; every rank holds the full matrix (1 M rows = 68 MB).
"""
from __future__ import annotations

import numpy as np

from . import capi


class ImportanceResult(dict):
    """``importances_mean``, ``importances_std``, ``importances`` like scikit-learn's Bunch, plus ``baseline_score``."""
    __getattr__ = dict.__getitem__


def shuffled_indices(n_samples: int, n_repeats: int, random_seed: int) -> np.ndarray:
    """Row r = where the permuted column's values come from at repeat r.  scikit-learn keeps shuffling one index vector
    and applies it to the *already permuted* column (``X_permuted[:, j] = X_permuted[shuffling_idx, j]``,
    SK/inspection/_permutation_importance.py:66-74), so the effective source index composes: e_r = e_{r-1}[s_r].
    numpy restatement (the checker of ``native_shuffled_indices``)."""
    rng = np.random.RandomState(random_seed)
    idx = np.arange(n_samples)
    eff = np.arange(n_samples)
    out = np.empty((n_repeats, n_samples), dtype=np.int32)
    for r in range(n_repeats):
        rng.shuffle(idx)
        eff = eff[idx]
        out[r] = eff
    return out


def native_shuffled_indices(n_samples: int, n_repeats: int, random_seed: int) -> np.ndarray:
    """The same vectors from ``hfe_perm_indices`` (numpy's legacy MT19937 stream and masked-rejection shuffle restated in
    C++, bit for bit): no Python-level arrays of 8-byte indices, half the time of five ``RandomState.shuffle`` calls."""
    import ctypes as C
    out = np.empty((n_repeats, n_samples), dtype=np.int32)
    capi.check(capi.load_library().hfe_perm_indices(int(random_seed) & 0xFFFFFFFF, n_samples, n_repeats,
                                                    out.ctypes.data_as(C.c_void_p)))
    return out


def shuffled_indices_device(n_samples: int, n_repeats: int, random_seed: int, device):
    """The same composition with the shuffles drawn on the device (``torch.randperm``): not scikit-learn's random
    stream, but the same procedure -- for cohorts where five host-side Fisher-Yates passes over the row index
    (~15 ms per million rows each, inherently serial) would dwarf the 86 scoring passes."""
    import torch
    gen = torch.Generator(device=device)
    gen.manual_seed(int(random_seed))
    eff = torch.arange(n_samples, device=device)
    out = torch.empty((n_repeats, n_samples), dtype=torch.int32, device=device)
    for r in range(n_repeats):
        eff = eff[torch.randperm(n_samples, generator=gen, device=device)]
        out[r] = eff.to(torch.int32)
    return out


def permutation_importance(clf, x, y, n_repeats: int = 5, random_state=None, rank: int = 0, world: int = 1,
                           shuffle: str = "sklearn"):
    """``x``: column-major float32 CUDA tensor ``[17, ld]`` (or a numpy ``[n,17]`` array), ``y``: class labels.
    ``shuffle="sklearn"`` reproduces scikit-learn's permutations exactly (host shuffles); ``shuffle="device"`` draws
    them on the GPU (same statistics, different random stream; identical on every rank for a given seed).
    Returns an ``ImportanceResult``; with ``world > 1`` torch.distributed must be initialised."""
    import torch

    if shuffle not in ("sklearn", "device"):
        raise ValueError("shuffle must be 'sklearn' or 'device'")

    if isinstance(x, np.ndarray):
        n = x.shape[0]
        x = clf.to_device_colmajor(x)
    else:
        n = int(np.asarray(y).shape[0])
    y_idx = np.searchsorted(clf.classes_, np.asarray(y, dtype=np.float64)).astype(np.uint8)
    dev = x.device
    y_dev = torch.from_numpy(y_idx).to(dev)
    rs = random_state if isinstance(random_state, np.random.RandomState) else np.random.RandomState(random_state)
    random_seed = rs.randint(np.iinfo(np.int32).max + 1)         # SK/inspection/_permutation_importance.py:278
    n_feat = clf.n_features_in_
    counts = torch.zeros(1 + n_feat * n_repeats, dtype=torch.int64, device=dev)
    status = clf.new_status()
    jobs = [(-1, 0)] + [(c, r) for c in range(n_feat) for r in range(n_repeats)]
    flags = torch.zeros(len(jobs), dtype=torch.int32, device=dev)      # status[0] of every pass (one array is reused)
    import ctypes as C
    lib, stream = capi.load_library(), torch.cuda.current_stream(dev).cuda_stream
    mine = list(range(len(jobs) * rank // world, len(jobs) * (rank + 1) // world))      # a contiguous block of passes per rank
    if shuffle == "sklearn":
        # one C call: host threads make scikit-learn's index vectors while the passes that need the earlier ones already
        # run on the device; this rank's passes in ascending repeat order, results scattered back to their slots
        order = sorted(mine, key=lambda k: (jobs[k][0] >= 0, jobs[k][1], jobs[k][0]))
        job_col = (C.c_int32 * len(order))(*[jobs[k][0] for k in order])
        job_rep = (C.c_int32 * len(order))(*[jobs[k][1] for k in order])
        perms_host = np.empty((n_repeats, n), dtype=np.int32)
        perms = torch.empty((n_repeats, n), dtype=torch.int32, device=dev)
        counts_run = torch.zeros(max(len(order), 1), dtype=torch.int64, device=dev)
        flags_run = torch.zeros(max(len(order), 1), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            capi.check(lib.hfe_permutation_passes(clf._handle, x.data_ptr(), n, x.stride(0), y_dev.data_ptr(),
                                                  int(random_seed) & 0xFFFFFFFF, n_repeats, len(order), job_col, job_rep,
                                                  perms_host.ctypes.data_as(C.c_void_p), perms.data_ptr(),
                                                  counts_run.data_ptr(), flags_run.data_ptr(), status.data_ptr(), stream))
        if order:
            at = torch.tensor(order, dtype=torch.long, device=dev)
            counts[at] = counts_run[:len(order)]
            flags[at] = flags_run[:len(order)]
    else:
        perms = shuffled_indices_device(n, n_repeats, random_seed, dev)
        # this rank's passes, column by column: one C call enqueues all repeats of a column back to back
        with torch.cuda.device(dev):
            at = 0
            while at < len(mine):
                col = jobs[mine[at]][0]
                run = [k for k in mine[at:] if jobs[k][0] == col]
                sel = (C.c_int32 * len(run))(*[jobs[k][1] for k in run])
                capi.check(lib.hfe_score_permuted(clf._handle, x.data_ptr(), n, x.stride(0), col,
                                                  perms.data_ptr() if col >= 0 else None, perms.stride(0), len(run), sel,
                                                  y_dev.data_ptr(), counts[run[0]:].data_ptr(), flags[run[0]:].data_ptr(),
                                                  status.data_ptr(), stream))
                at += len(run)
    flags_host = flags.cpu().numpy()
    if (flags_host & capi.ROW_NONFINITE).any():
        raise ValueError("Input X contains NaN or infinity.")          # what scikit-learn's scorer raises
    for k in np.nonzero(flags_host & capi.WORKLIST_OVERFLOW)[0]:
        # more rows than the work list holds needed float64 in this pass: the count is short by the rows beyond the
        # capacity, so recount the pass from the all-float64 labels of the materialised shuffle
        col, rep = jobs[k]
        xp = x
        if col >= 0:
            xp = x.clone()
            xp[col, :n] = x[col, perms[rep].long()]
        lab = torch.empty(n, dtype=torch.uint8, device=dev)
        clf.score_f64_device(xp, n=n, labels=lab)
        counts[k] = (lab == y_dev).sum()
    if world > 1:
        torch.distributed.all_reduce(counts)
    acc = counts.cpu().numpy().astype(np.float64) / n
    baseline = acc[0]
    scores = acc[1:].reshape(n_feat, n_repeats)
    importances = baseline - scores
    return ImportanceResult(importances_mean=importances.mean(axis=1), importances_std=importances.std(axis=1),
                            importances=importances, baseline_score=baseline)
