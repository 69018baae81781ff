"""BASELINE config 4 (2000 trees, depth <= 16): the three ways of walking the forest side by side (development aid)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
from oracle import stack_oracle as so  # noqa: E402  (checker)
n = int(os.environ.get("STRESS_ROWS", 2_000_000))
base = P.export.read_blob(open(os.path.join(ROOT, "tests/golden/published_model.blob"), "rb").read())
arrays = P.synth_models.stress_stack(base, with_svc=False)
clf = P.HFStackB200.from_arrays(arrays)
x = P.synth_rows_device(P.synth.schema_specs(), 2020, 0, n)
idx = np.arange(0, n, n // 400)[:400]
Xs = x[:, torch.from_numpy(idx).cuda()].T.cpu().numpy().astype(np.float64)
want = so.predict_proba(arrays, Xs)
ref = None
for eng in (sys.argv[1:] or ["l1", "l1+l2window", "smem"]):
    clf.set_tree_engine(eng)
    out = torch.empty((n, 2), dtype=torch.float32, device="cuda"); st = clf.new_status()
    for _ in range(2): clf.predict_proba_device(x, n=n, out=out, status=st)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(5)]
    for a, b in ev:
        a.record(); clf.predict_proba_device(x, n=n, out=out, status=st); b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)[2]
    err = np.abs(out[torch.from_numpy(idx).cuda()].cpu().numpy() - want).max()
    same = "" if ref is None else f" identical to l1: {bool(torch.equal(out, ref))}"
    if ref is None: ref = out.clone()
    print(f"{eng:12s} n={n} {ms:.2f} ms {n/ms/1e3:.2f} M rows/s  {n*2000/ms/1e6:.1f} G tree walks/s  max err vs oracle {err:.2e} fix={int(st[1])} status={int(st[0])}{same}", flush=True)
