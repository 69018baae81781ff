"""Per-call time of the streaming pass: through the facade vs straight through the C ABI (development aid)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
n = 10_000_000
arrays = P.export.read_blob(open(os.path.join(ROOT, "tests/golden/retrained_model.blob"), "rb").read())
clf = P.HFStackB200.from_arrays(P.synth_models.without_member(arrays, "svc"))
cand = P.synth.candidate_specs()
specs = [cand[c] for c in np.load(os.path.join(ROOT, "tests/golden/retrained_columns.npy"))]
x = P.synth_rows_device(specs, 2020, 0, n)
out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
st = clf.new_status()
lib = clf._lib
stream = torch.cuda.current_stream().cuda_stream
def facade(): clf.predict_proba_device(x, n=n, out=out, status=st)
def direct(): lib.hfe_predict_proba(clf._handle, x.data_ptr(), n, x.stride(0), out.data_ptr(), st.data_ptr(), stream)
for name, fn in (("facade", facade), ("direct C ABI", direct), ("facade", facade), ("direct C ABI", direct)):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50): fn()
    b.record(); torch.cuda.synchronize()
    print(f"{name:14s} {a.elapsed_time(b) / 50 * 1e3:.1f} us per call = {n / (a.elapsed_time(b) / 50) / 1e6:.1f} G rows/s")
