"""Time the fused kernel (development aid): python tools/time_kernel.py [model] [rows]; honours HFE_LIB."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "published"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
gold = os.path.join(ROOT, "tests", "golden")
clf = P.HFStackB200.from_blob_file(os.path.join(gold, f"{name}_model.blob"))
specs = P.synth.schema_specs() if name == "published" else [P.synth.candidate_specs()[c] for c in np.load(os.path.join(gold, "retrained_columns.npy"))]
x = P.synth_rows_device(specs, 2020, 0, n)
out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
st = clf.new_status()
for _ in range(3):
    clf.predict_proba_device(x, n=n, out=out, status=st)
torch.cuda.synchronize()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
for a, b in ev:
    a.record(); clf.predict_proba_device(x, n=n, out=out, status=st); b.record()
torch.cuda.synchronize()
ms = sorted(a.elapsed_time(b) for a, b in ev)[5]
err_rows = min(n, int(os.environ.get("ERR_ROWS", 500_000)))
o64 = clf.predict_proba_f64_device(x, n=err_rows)
err = float((out[:err_rows].double() - o64).abs().max().item())
print(f"{os.environ.get('HFE_LIB', 'default'):40s} {name} n={n} {ms:.3f} ms  {n / ms / 1e6:.3f} G rows/s  fix={int(st[1].item())} maxerr_vs_f64={err:.2e} (first {err_rows} rows)")
