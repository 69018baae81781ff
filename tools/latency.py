"""Single-patient latency of the host call (predict_hf.py:36) next to scikit-learn's on the same rebuilt model."""
import os, sys, time, warnings
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
from oracle import sk_rebuild  # noqa: E402  (baseline only)
warnings.simplefilter("ignore")
blob = os.path.join(ROOT, "tests/golden/published_model.blob")
clf = P.HFStackB200.from_blob_file(blob)
sk = sk_rebuild.rebuild_stack(P.export.read_blob(open(blob, "rb").read()))
x = np.reshape([1, 1, 0, 0, 1, 0, 1, 1, 0, 0, 0, 0, 0, 13, 0, 0, 55], (1, -1))
for name, f in (("b200 facade", clf.predict_proba), ("scikit-learn", sk.predict_proba)):
    for _ in range(20): f(x)
    t = []
    for _ in range(200):
        t0 = time.perf_counter(); f(x); t.append(time.perf_counter() - t0)
    t.sort()
    print(f"{name:14s} 1 patient: median {1e6 * t[100]:.1f} us, p95 {1e6 * t[190]:.1f} us  -> {100 * f(x)[0, 1]:.2f} %")

X = P.synth.host_rows(P.synth.schema_specs(), 1, 0, 10_000).astype(np.float64)      # BASELINE config 1: 10 k patients
for name, f in (("b200 facade", clf.predict_proba), ("scikit-learn", sk.predict_proba)):
    for _ in range(3): f(X)
    t = []
    for _ in range(20):
        t0 = time.perf_counter(); f(X); t.append(time.perf_counter() - t0)
    t.sort()
    print(f"{name:14s} 10 000 patients: median {1e3 * t[10]:.3f} ms")
