"""Where the warps of the tensor-core kernel wait (development aid; needs a library built with -DHFE_MMA_TRACE)."""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
name = sys.argv[1] if len(sys.argv) > 1 else "published"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
gold = os.path.join(ROOT, "tests", "golden")
clf = P.HFStackB200.from_blob_file(os.path.join(gold, f"{name}_model.blob"))
specs = P.synth.schema_specs()
x = P.synth_rows_device(specs, 2020, 0, n)
out = torch.empty((n, 2), dtype=torch.float32, device="cuda")
st = clf.new_status()
lib = P.capi.load_library()
clf.set_rbf_engine("tensor")
for _ in range(2):
    clf.predict_proba_device(x, n=n, out=out, status=st)
torch.cuda.synchronize()
buf = (C.c_ulonglong * 32)()
lib.hfe_debug_trace(buf, 1)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); clf.predict_proba_device(x, n=n, out=out, status=st); b.record()
torch.cuda.synchronize()
lib.hfe_debug_trace(buf, 0)
t = np.array(list(buf), dtype=np.float64)
tiles = (n + 127) // 128
names = {0: "acc_full blk0", 1: "acc_full blk1", 2: "acc_full blk2", 3: "acc_full blk3", 4: "acc_full blk4", 8: "issuer a_full", 9: "issuer acc_empty", 10: "issuer issue_next (incl a_full)",
         11: "producer a_empty", 12: "producer ring_free", 13: "finisher sum_full", 14: "exponent loop total"}
print(f"{a.elapsed_time(b):.3f} ms, {tiles} tiles; cycles per tile per waiting warp:")
div = {0: 8, 1: 8, 2: 8, 3: 8, 4: 8, 8: 2, 9: 2, 10: 2, 11: 4, 12: 4, 13: 4, 14: 8}
for k, v in names.items():
    print(f"  {v:34s} {t[k] / tiles / div[k] * (2 if k < 11 or k == 14 else 1):10.1f}")
print("  (exponent / issuer figures are per tile OF THE WARP'S STREAM)")
