"""Stress forest on rows in their natural order vs the same rows sorted by feature tuple (how much would coherent lanes buy?)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
n = 2_000_000
base = P.export.read_blob(open(os.path.join(ROOT, "tests/golden/published_model.blob"), "rb").read())
clf = P.HFStackB200.from_arrays(P.synth_models.stress_stack(base, with_svc=False))
x = P.synth_rows_device(P.synth.schema_specs(), 2020, 0, n)
key = torch.zeros(n, dtype=torch.float64, device="cuda")
for f in range(17):
    key = key * 64.0 + x[f, :n].double()          # lexicographic key (values are small integers)
order = torch.argsort(key)
xs = x.clone(); xs[:, :n] = x[:, order]
# block-local variant: sort within consecutive groups of 2048 rows only
grp = torch.arange(n, device="cuda") // 2048
order2 = torch.argsort(grp.double() * 1e30 + key) if False else torch.cat([order_g + s for s, order_g in
          ((s, torch.argsort(key[s:s + 2048])) for s in range(0, n, 2048))])
xb = x.clone(); xb[:, :n] = x[:, order2]
for name, xx in (("natural", x), ("globally sorted", xs), ("sorted within 2048-row blocks", xb)):
    out = torch.empty((n, 2), dtype=torch.float32, device="cuda"); st = clf.new_status()
    for _ in range(2): clf.predict_proba_device(xx, n=n, out=out, status=st)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(5)]
    for a, b in ev:
        a.record(); clf.predict_proba_device(xx, n=n, out=out, status=st); b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)[2]
    print(f"{name:32s} {ms:.2f} ms {n / ms / 1e3:.1f} M rows/s", flush=True)
