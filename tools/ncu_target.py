"""Small single-GPU target for ncu: N warm launches of the fused kernel on synthetic rows (NCU_ROWS, NCU_MODEL)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402

name = os.environ.get("NCU_MODEL", "published")
n = int(os.environ.get("NCU_ROWS", 4_000_000))
gold = os.path.join(ROOT, "tests", "golden")
base = "retrained" if name == "retrained" else "published"
arrays = P.export.read_blob(open(os.path.join(gold, f"{base}_model.blob"), "rb").read())
if name == "nosvc":
    arrays = P.synth_models.without_member(arrays, "svc")
elif name == "stress":
    arrays = P.synth_models.stress_stack(arrays, with_svc=False)
clf = P.HFStackB200.from_arrays(arrays)
if base == "published":
    specs = P.synth.schema_specs()
else:
    cand = P.synth.candidate_specs()
    specs = [cand[c] for c in np.load(os.path.join(gold, "retrained_columns.npy"))]
x = P.synth_rows_device(specs, 2020, 0, n)
out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
st = clf.new_status()
for _ in range(int(os.environ.get("NCU_LAUNCHES", 4))):
    clf.predict_proba_device(x, n=n, out=out, status=st)
torch.cuda.synchronize()
print("fixups", int(st[1].item()), "p1[0]", float(out[0, 1].item()))
