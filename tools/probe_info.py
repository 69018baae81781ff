"""Print what the creation-time guard probe saw for the two fixture models (hfe_model_info)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402

gold = os.path.join(ROOT, "tests", "golden")
for name in ("published", "retrained"):
    clf = P.HFStackB200(open(os.path.join(gold, f"{name}_model.blob"), "rb").read(), device=0)
    for scale in (1.0, 1e-3):
        clf.calibrate_guards(scale)
        i = clf.info
        print(name, "scale", scale, "svc_guard %.3e observed %.3e label_guard %.3e observed %.3e widened %d rows %d simt %.3e tensor %.3e" % (
            i.svc_guard, i.svc_observed_err, i.label_guard, i.label_observed_err, i.guard_widened, i.probe_rows, i.svc_observed_by_engine[0], i.svc_observed_by_engine[1]))
