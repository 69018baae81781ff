"""Soak test of the tensor-core engine: random row counts, both fixture models, every call compared with the SIMT engine
(labels equal, probabilities within 5e-7) -- looks for rare protocol faults (a stuck barrier traps after ~2 s)."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
rng = np.random.default_rng(int(os.environ.get("SOAK_SEED", 1)))
gold = os.path.join(ROOT, "tests", "golden")
cand = P.synth.candidate_specs()
models = {"published": P.synth.schema_specs(), "retrained": [cand[c] for c in np.load(os.path.join(gold, "retrained_columns.npy"))]}
t_end = time.time() + float(os.environ.get("SOAK_SECONDS", 40))
calls = 0
for name, specs in models.items():
    clf = P.HFStackB200.from_blob_file(os.path.join(gold, f"{name}_model.blob"))
    xbig = P.synth_rows_device(specs, 7, 0, 3_000_000)
    while time.time() < t_end - (20 if name == "published" else 0):
        n = int(rng.choice([rng.integers(1, 300), rng.integers(1, 40000), rng.integers(1, 3_000_000)]))
        off = int(rng.integers(0, (3_000_000 - n) // 4 + 1)) * 4
        x = xbig[:, off:]
        outs = {}
        for eng in ("simt", "tensor"):
            clf.set_rbf_engine(eng)
            out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
            lab = torch.empty(n, dtype=torch.uint8, device="cuda")
            clf.score_device(x, n=n, out=out, labels=lab)
            outs[eng] = (out, lab)
        torch.cuda.synchronize()
        assert torch.equal(outs["simt"][1], outs["tensor"][1]), (name, n, off)
        assert float((outs["simt"][0] - outs["tensor"][0]).abs().max()) <= 5e-7, (name, n, off)
        calls += 1
    clf.close()
print(f"soak ok: {calls} random-size calls per engine, labels equal, probabilities within 5e-7")
