"""Small run of every kernel for compute-sanitizer (memcheck): ragged sizes, permuted column, general trees, no-alloc."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402

gold = os.path.join(ROOT, "tests", "golden")
base = P.export.read_blob(open(os.path.join(gold, "published_model.blob"), "rb").read())
specs = P.synth.schema_specs()
models = [base, P.synth_models.without_member(base, "svc"),
          P.synth_models.stress_stack(base, with_svc=True, n_trees=24, n_leaves=64),
          P.synth_models.stress_stack(base, with_svc=False, n_trees=24, n_leaves=64)]
for arrays in models:
    clf = P.HFStackB200.from_arrays(arrays)
    for n in (1, 127, 1029):
        x = P.synth_rows_device(specs, 5, 0, n)
        out = clf.predict_proba_device(x, n=n)
        o64 = clf.predict_proba_f64_device(x, n=n)
        lab = torch.empty(n, dtype=torch.uint8, device="cuda")
        perm = torch.from_numpy(np.random.RandomState(0).permutation(n).astype(np.int32)).cuda()
        score = torch.zeros(1, dtype=torch.int64, device="cuda")
        clf.score_device(x, n=n, perm_col=13, perm_idx=perm, out=out, labels=lab, y=lab, score=score)
        clf.apply_device(x, n=n)
        P.metrics.device_counts(out, lab)
        torch.cuda.synchronize()
        assert float((out.double() - o64).abs().max()) < 1.0
    X = P.synth.host_rows(specs, 6, 0, 333).astype(np.float64)
    clf.predict_proba(X); clf.predict(X); clf.transform(X)
    clf.predict_proba_pinned(torch.from_numpy(X.astype(np.float32)).pin_memory(), chunk_rows=128)
    lib = P.capi.load_library()
    x = clf.to_device_colmajor(X)
    o = torch.empty((333, 2), dtype=torch.float32, device="cuda")
    P.capi.check(lib.hfe_predict_proba_noalloc(clf._handle, x.data_ptr(), 333, x.stride(0), o.data_ptr(), None, None))
    torch.cuda.synchronize()
    clf.close()
print("sanitize target done")
