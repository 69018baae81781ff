"""Model-handle creation time (blob -> tables on the device, incl. the guard probe) and first-call latency."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
blob = open(os.path.join(ROOT, "tests/golden/published_model.blob"), "rb").read()
import torch
torch.cuda.init(); torch.zeros(1, device="cuda")
for k in range(4):
    t0 = time.perf_counter(); clf = P.HFStackB200(blob, device=0); t1 = time.perf_counter()
    p = clf.predict_proba(np.array([[1, 1, 0, 0, 1, 0, 1, 1, 0, 0, 0, 0, 0, 13, 0, 0, 55]])); t2 = time.perf_counter()
    print(f"create {1e3 * (t1 - t0):.2f} ms, first predict {1e3 * (t2 - t1):.3f} ms, p1 {p[0, 1]:.6f}")
    clf.close()
