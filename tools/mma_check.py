"""Development check of the tensor-core RBF engine against the float64 kernel and the SIMT engine (run under gpurun)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "published"
sizes = [int(a) for a in sys.argv[2:]] or [128, 1000, 100_000, 10_000_000]
gold = os.path.join(ROOT, "tests", "golden")
clf = P.HFStackB200.from_blob_file(os.path.join(gold, f"{name}_model.blob"))
if name == "published":
    specs = P.synth.schema_specs()
else:
    cand = P.synth.candidate_specs()
    specs = [cand[c] for c in np.load(os.path.join(gold, "retrained_columns.npy"))]
print("engine (auto):", clf.info.rbf_engine, "n_sv", clf.info.n_support_vectors, flush=True)
for n in sizes:
    x = P.synth_rows_device(specs, 2020, 0, n)
    res = {}
    for eng in ("simt", "tensor"):
        clf.set_rbf_engine(eng)
        out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
        dec = torch.empty(n, dtype=torch.float64, device="cuda")
        lab = torch.empty(n, dtype=torch.uint8, device="cuda")
        st = clf.new_status()
        clf.score_device(x, n=n, out=out, labels=lab, decision=dec, status=st)
        torch.cuda.synchronize()
        t = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            clf.predict_proba_device(x, n=n, out=out, status=st)
            b.record()
            torch.cuda.synchronize()
            t.append(a.elapsed_time(b))
        res[eng] = (out.clone(), dec.clone(), lab.clone(), st[:4].cpu().tolist(), sorted(t)[2])
    m = min(n, 2_000_000)
    o64 = torch.empty((m, 2), dtype=torch.float64, device="cuda")
    d64 = torch.empty(m, dtype=torch.float64, device="cuda")
    l64 = torch.empty(m, dtype=torch.uint8, device="cuda")
    clf.score_f64_device(x, n=m, out=o64, labels=l64, decision=d64)
    torch.cuda.synchronize()
    for eng in ("simt", "tensor"):
        out, dec, lab, st, ms = res[eng]
        print(f"n={n:>9} {eng:>6}: {ms:8.3f} ms = {n / ms / 1e6:8.3f} G rows/s  status={st}  "
              f"max|p-p64|={float((out[:m].double() - o64).abs().max()):.3e}  max|z-z64|={float((dec[:m] - d64).abs().max()):.3e}  "
              f"labels_equal={bool(torch.equal(lab[:m], l64))}", flush=True)
    print(f"             simt vs tensor: max|dp|={float((res['simt'][0] - res['tensor'][0]).abs().max()):.3e}", flush=True)
# values that one bfloat16 cannot carry: the extension block
n = 200_000
X = P.synth.host_rows(specs, 5, 0, n).astype(np.float64)
rng = np.random.default_rng(0)
X += rng.normal(0, 0.37, X.shape) * (rng.random(X.shape) < 0.3)
x = clf.to_device_colmajor(X)
o64 = torch.empty((n, 2), dtype=torch.float64, device="cuda")
clf.score_f64_device(x, n=n, out=o64)
for eng in ("simt", "tensor"):
    clf.set_rbf_engine(eng)
    out = clf.predict_proba_device(x, n=n)
    torch.cuda.synchronize()
    print(f"fractional inputs {eng:>6}: max|p-p64|={float((out.double() - o64).abs().max()):.3e}", flush=True)
print("done")
