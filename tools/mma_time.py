"""Time one scoring call per engine on 10 M rows (development aid); honours HFE_LIB."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
name = sys.argv[1] if len(sys.argv) > 1 else "published"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
engines = sys.argv[3].split(",") if len(sys.argv) > 3 else ["simt", "tensor"]
gold = os.path.join(ROOT, "tests", "golden")
clf = P.HFStackB200.from_blob_file(os.path.join(gold, f"{name}_model.blob"))
if name == "published":
    specs = P.synth.schema_specs()
else:
    cand = P.synth.candidate_specs()
    specs = [cand[c] for c in np.load(os.path.join(gold, "retrained_columns.npy"))]
x = P.synth_rows_device(specs, 2020, 0, n)
out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
st = clf.new_status()
for eng in engines:
    clf.set_rbf_engine(eng)
    for _ in range(3):
        clf.predict_proba_device(x, n=n, out=out, status=st)
    torch.cuda.synchronize()
    t = []
    for _ in range(7):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); clf.predict_proba_device(x, n=n, out=out, status=st); b.record()
        torch.cuda.synchronize()
        t.append(a.elapsed_time(b))
    ms = sorted(t)[3]
    print(f"{os.environ.get('HFE_LIB', 'default'):34s} {name} {eng:>6} n={n} {ms:.3f} ms {n / ms / 1e6:.3f} G rows/s fix={int(st[1])}", flush=True)
