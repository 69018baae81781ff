"""hfe_model_plan: the host-only twin of hfe_model_create -- blob validation and table planning without a GPU."""
import copy
import ctypes as C

import numpy as np
import pytest


def plan(pkg, lib, arrays_or_blob):
    blob = arrays_or_blob if isinstance(arrays_or_blob, (bytes, bytearray)) else pkg.export.write_blob(arrays_or_blob)
    info = pkg.capi.ModelInfo()
    order = (C.c_int32 * 17)()
    rc = lib.hfe_model_plan(bytes(blob), len(blob), C.byref(info), order)
    return rc, info, list(order), lib.hfe_last_error_string().decode()


def test_plan_of_the_fixture_models(pkg, cuda_lib, published, retrained):
    for fx in (published, retrained):
        rc, info, order, _ = plan(pkg, cuda_lib, fx["blob"])
        assert rc == 0
        svc = fx["arrays"].members[0]
        n_sv = svc.sv.shape[0]
        assert info.n_support_vectors == n_sv and info.n_trees == 100 and info.tree_mode == 1
        assert sorted(order) == list(range(17))
        # the 9 leading internal features are the ones with the fewest distinct prefixes among the support vectors:
        # re-derive the greedy choice here
        chosen, rest = [], list(range(17))
        for _ in range(9):
            best = min(rest, key=lambda c: (len(np.unique(svc.sv[:, chosen + [c]], axis=0)), c))
            chosen.append(best)
            rest.remove(best)
        assert order[:9] == chosen and order[9:] == rest
        flat = (n_sv + 3) // 4 * 4 * 18
        assert info.rbf_fma_per_row % 36 == 0 and 0 < info.rbf_fma_per_row < 0.75 * flat
        assert info.device_bytes > n_sv * 17 * 8


def test_plan_refuses_what_model_create_refuses(pkg, cuda_lib, published):
    base = published["arrays"]

    def broken(mutate):
        arrays = copy.deepcopy(base)
        mutate(arrays)
        return plan(pkg, cuda_lib, arrays)

    def bad_n_support(a):
        a.members[0].n_support = np.array([1, 2])                # SK/svm/_base.py:680-686 consistency check

    def nan_vector(a):
        a.members[0].sv[3, 4] = np.nan

    def zero_scale(a):
        a.members[0].scale[2] = 0.0

    def child_before_parent(a):
        t = a.members[1].trees[5]
        t.left[0], t.right[0] = 0, 2

    def feature_out_of_range(a):
        a.members[1].trees[7].feature[0] = 17

    def nonfinite_leaf(a):
        a.members[1].trees[1].value[1] = np.inf

    def nonfinite_combiner(a):
        a.final_coef = np.array([1.0, np.nan, 2.0])

    for mutate, needle in [(bad_n_support, "n_support"), (nan_vector, "non-finite"), (zero_scale, "scale_"),
                           (child_before_parent, "child"), (feature_out_of_range, "feature"),
                           (nonfinite_leaf, "leaf"), (nonfinite_combiner, "combiner")]:
        try:
            rc, _, _, msg = broken(mutate)
        except Exception as exc:                                   # the Python writer may already refuse it
            assert isinstance(exc, (ValueError, pkg.export.walker.ModelShapeError)), mutate.__name__
            continue
        assert rc in (2, 3), (mutate.__name__, rc)                 # HFE_ERR_BAD_BLOB / HFE_ERR_UNSUPPORTED
        assert needle in msg, (mutate.__name__, msg)
    rc, _, _, msg = plan(pkg, cuda_lib, b"\0" * 200)
    assert rc == 2 and "magic" in msg


def test_plan_of_stacks_without_svc_keeps_the_feature_order(pkg, cuda_lib, published):
    no_svc = pkg.synth_models.without_member(published["arrays"], "svc")
    rc, info, order, _ = plan(pkg, cuda_lib, no_svc)
    assert rc == 0 and info.n_support_vectors == 0 and info.rbf_fma_per_row == 0
    assert order == list(range(17))
    rng = np.random.default_rng(5)
    scattered = copy.deepcopy(published["arrays"])
    scattered.members[0].sv = scattered.members[0].sv + rng.normal(0, 0.05, scattered.members[0].sv.shape)
    rc, info, _, _ = plan(pkg, cuda_lib, scattered)
    assert rc == 0 and info.rbf_fma_per_row == (434 + 3) // 4 * 4 * 18      # nothing shared: flat layout
