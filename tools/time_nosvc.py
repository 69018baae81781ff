"""Time the SVC-free (HBM-bound) configuration of the fused kernel (development aid); honours HFE_LIB."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
n = 10_000_000
arrays = P.synth_models.without_member(P.export.read_blob(open(os.path.join(ROOT, "tests/golden/published_model.blob"), "rb").read()), "svc")
clf = P.HFStackB200.from_arrays(arrays)
x = P.synth_rows_device(P.synth.schema_specs(), 2020, 0, n)
out = torch.empty((n, 2), dtype=torch.float32, device="cuda"); st = clf.new_status()
for _ in range(5): clf.predict_proba_device(x, n=n, out=out, status=st)
torch.cuda.synchronize()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
for a, b in ev:
    a.record(); clf.predict_proba_device(x, n=n, out=out, status=st); b.record()
torch.cuda.synchronize()
ms = sorted(a.elapsed_time(b) for a, b in ev)[10]
o64 = clf.predict_proba_f64_device(x, n=n)
err = float((out.double() - o64).abs().max().item())
print(f"{os.environ.get('HFE_LIB','default'):36s} nosvc {ms:.4f} ms {n/ms/1e6:.2f} G rows/s  hbm frac {n*76/ms/1e6/6545.6:.3f}  "
      f"fix={int(st[1].item())} maxerr_vs_f64={err:.2e} ({n} rows)")
