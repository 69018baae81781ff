"""Time of scikit-learn's five cumulative shuffles of 1 M indices: numpy vs hfe_perm_indices (host only)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
ins = P.inspection
for _ in range(3):
    t0 = time.perf_counter(); a = ins.native_shuffled_indices(1_000_000, 5, 7); t1 = time.perf_counter()
    print(f"native {1e3 * (t1 - t0):.1f} ms")
t0 = time.perf_counter(); b = ins.shuffled_indices(1_000_000, 5, 7); t1 = time.perf_counter()
print(f"numpy {1e3 * (t1 - t0):.1f} ms, equal {np.array_equal(a, b)}")
