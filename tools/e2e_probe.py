"""Host-call throughput (clf.predict_proba on pageable float64 rows) against the number of host threads."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402

n = int(os.environ.get("E2E_ROWS", 10_000_000))
blob = open(os.path.join(ROOT, "tests", "golden", "retrained_model.blob"), "rb").read()
print("cpus available", len(os.sched_getaffinity(0)), "of", os.cpu_count())
os.system("lscpu | egrep 'Model name|Socket|NUMA node|^CPU\\(s\\)' ; numactl -H 2>/dev/null | head -5")
cand = P.synth.candidate_specs()
specs = [cand[c] for c in np.load(os.path.join(ROOT, "tests", "golden", "retrained_columns.npy"))]
X = P.synth_rows_device(specs, 2020, 0, n)[:, :n].T.contiguous().cpu().numpy().astype(np.float64)
for threads in sys.argv[1:] or ["8", "16", "32"]:
    os.environ["HFE_HOST_THREADS"] = threads
    clf = P.HFStackB200(blob, device=0)
    clf.predict_proba(X[:1_000_000])
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        clf.predict_proba(X)
        best = min(best, time.perf_counter() - t0)
    print(f"threads {threads}: best {best * 1e3:.1f} ms = {n / best / 1e6:.0f} M rows/s")
    clf.close()
# plain host copy bandwidth for scale: numpy copy of the same matrix (read 1.36 GB + write 1.36 GB, one thread)
t0 = time.perf_counter(); Y = X.copy(); t1 = time.perf_counter()
print(f"numpy copy 1 thread: {2 * X.nbytes / (t1 - t0) / 1e9:.1f} GB/s")
