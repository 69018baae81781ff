"""BASELINE.json config 4 on the configuration itself: the 2000-tree, depth <= 16 stress forest (2.05 M nodes, node
tables far beyond shared memory) with and without the SVC member, through the C ABI.

Reference arithmetic: the tree-major loop of SK/ensemble/_gradient_boosting.pyx:25-73 over ``Node`` records
(SK/tree/_tree.pxd:15-25) and ``Tree.apply`` (SK/tree/_tree.pyx:954-996).  The oracle is first pinned at this scale
against scikit-learn's own forest built through ``Tree.__setstate__`` (oracle/sk_rebuild.py), then the CUDA path is
compared with the oracle: probabilities <= 1e-6, labels and leaf ids bit-exact.
"""
import warnings

import numpy as np
import pytest
from scipy.special import expit

from oracle import sk_rebuild
from oracle import stack_oracle as so

PROBA_TOL = 1e-6


@pytest.fixture(scope="module")
def stress(pkg, published):
    full = pkg.synth_models.stress_stack(published["arrays"], with_svc=True)       # ~10 s of Python tree building
    return {True: full, False: pkg.synth_models.without_member(full, "svc")}


def test_oracle_is_pinned_to_sklearn_on_the_stress_forest(pkg, published, stress):
    arrays = stress[False]
    gbt = next(m for m in arrays.members if m.kind == "gbt")
    assert len(gbt.trees) == 2000 and sum(t.left.shape[0] for t in gbt.trees) == 2000 * 1023
    X = pkg.synth.host_rows(published["specs"], 77, 0, 300).astype(np.float64)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        stack = sk_rebuild.rebuild_stack(arrays)
        want = stack.predict_proba(X)
        leaves = stack.estimators_[0].apply(X)[:, :, 0].astype(np.int64)
    assert np.abs(so.predict_proba(arrays, X) - want).max() <= 1e-12
    assert np.array_equal(so.gbt_apply(gbt, X), leaves)


@pytest.mark.gpu
@pytest.mark.parametrize("with_svc", [False, True])
def test_stress_forest_parity(pkg, cuda_lib, published, stress, with_svc):
    import torch
    arrays = stress[with_svc]
    clf = pkg.HFStackB200.from_arrays(arrays)
    info = clf.info
    assert info.tree_mode == 2 and info.n_trees == 2000 and info.n_tree_nodes == 2000 * 1023
    assert info.max_tree_depth <= 16 and info.max_tree_depth >= 12
    # 2 M rows on the device: error bits clean, float64 re-evaluations far below the work-list capacity (the boosted
    # sum's guard band grows with 2000 trees), output a probability pair
    n = 2_000_000
    x = pkg.synth_rows_device(published["specs"], 2020, 0, n)
    st = clf.new_status()
    out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
    lab = torch.empty(n, dtype=torch.uint8, device="cuda")
    clf.score_device(x, n=n, out=out, labels=lab, status=st)
    torch.cuda.synchronize()
    assert int(st[0].item()) == 0
    assert int(st[2].item()) < pkg.capi.WORKLIST_CAPACITY // 4, int(st[2].item())
    assert bool(((out[:, 0] + out[:, 1]) - 1.0).abs().max() <= 1.2e-7)
    # oracle on >= 20 000 rows spread over the batch
    idx = np.arange(0, n, n // 20000)[:20000]
    sel = torch.from_numpy(idx).cuda()
    Xs = x[:, sel].T.cpu().numpy().astype(np.float64)
    z = so.decision_function(arrays, Xs)
    p1 = expit(z)
    got = out[sel].cpu().numpy().astype(np.float64)
    assert np.abs(got - np.stack([1.0 - p1, p1], axis=1)).max() <= PROBA_TOL
    assert np.array_equal(lab[sel].cpu().numpy(), (z > 0).astype(np.uint8))
    # leaf ids (GradientBoostingClassifier.apply) bit-exact on 2 000 rows x 2 000 trees
    gbt = next(m for m in arrays.members if m.kind == "gbt")
    leaves = clf.apply_device(x[:, sel[:2000]].contiguous(), n=2000).cpu().numpy().view(np.uint16).astype(np.int64)
    assert np.array_equal(leaves, so.gbt_apply(gbt, Xs[:2000]))
    # labels of the fast pass == labels of the all-float64 pass on every one of the 2 M rows
    lab64 = torch.empty(n, dtype=torch.uint8, device="cuda")
    clf.score_f64_device(x, n=n, labels=lab64)
    torch.cuda.synchronize()
    assert torch.equal(lab, lab64)
    clf.close()
