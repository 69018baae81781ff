"""BASELINE.json config 5: permutation importance over 1 M rows -- 17 columns x 5 repeats + baseline = 86 passes.
Run plainly (1 GPU) or under torch.distributed.run (passes dealt round-robin over the ranks)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402

rank, local_rank, world = P.shard.init_process_group()
torch.cuda.set_device(local_rank)
n = int(os.environ.get("PERM_ROWS", 1_000_000))
gold = os.path.join(ROOT, "tests", "golden")
clf = P.HFStackB200.from_blob_file(os.path.join(gold, "published_model.blob"), device=local_rank)
specs = P.synth.schema_specs()
x = P.synth_rows_device(specs, 2020, 0, n, device=local_rank)
lab = torch.empty(n, dtype=torch.uint8, device=x.device)
clf.score_device(x, n=n, labels=lab)
flip = torch.from_numpy((P.synth.hash_u32(99, np.arange(n), 251) < 2 ** 29)).to(x.device)     # 12.5 % label noise
y = (lab ^ flip.to(torch.uint8)).cpu().numpy().astype(np.float64)
P.inspection.permutation_importance(clf, x, y, n_repeats=1, random_state=0, rank=rank, world=world)  # warm-up
torch.cuda.synchronize()
for mode in ("sklearn", "device"):
    P.inspection.permutation_importance(clf, x, y, n_repeats=1, random_state=0, rank=rank, world=world, shuffle=mode)
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    t0 = time.perf_counter()
    res = P.inspection.permutation_importance(clf, x, y, n_repeats=5, random_state=2020, rank=rank, world=world,
                                              shuffle=mode)
    torch.cuda.synchronize()
    dt = P.shard.max_over_ranks(time.perf_counter() - t0, world, device=x.device)
    if rank == 0:
        order = np.argsort(-res.importances_mean)[:5]
        print(json.dumps({"workload": "permutation importance, published stack", "shuffle": mode, "rows": n, "passes": 86,
                          "n_gpus": world, "seconds": dt, "rows_scored_per_s": 86 * n / dt,
                          "baseline_accuracy": res.baseline_score,
                          "top_features": [(P.synth.SCHEMA_NAMES[i], round(float(res.importances_mean[i]), 5))
                                           for i in order]}))
if world > 1:
    torch.distributed.destroy_process_group()
