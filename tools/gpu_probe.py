"""Ad-hoc GPU probe used during development (run under gpurun): parity on the golden rows, float32-vs-float64 path
agreement at scale, and a first timing of the fused kernel.  Not part of the test suite."""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402
from oracle import stack_oracle as so  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def specs_for(name):
    if name == "published":
        return P.synth.schema_specs()
    cols = np.load(os.path.join(GOLD, "retrained_columns.npy"))
    cand = P.synth.candidate_specs()
    return [cand[c] for c in cols]


def main():
    n_big = int(os.environ.get("PROBE_ROWS", 10_000_000))
    out = {}
    for name in ("published", "retrained"):
        clf = P.HFStackB200.from_blob_file(os.path.join(GOLD, f"{name}_model.blob"))
        model = P.export.read_blob(open(os.path.join(GOLD, f"{name}_model.blob"), "rb").read())
        g = np.load(os.path.join(GOLD, f"golden_{name}.npz"))
        X = g["X"].astype(np.float64)
        proba = clf.predict_proba(X)
        res = {
            "golden_rows": int(X.shape[0]),
            "max_dp": float(np.abs(proba - g["proba"]).max()),
            "labels_exact": bool(np.array_equal(clf.predict(X), g["label"])),
            "leaves_exact": bool(np.array_equal(clf.apply(X), g["leaves"].astype(np.int64))),
            "max_dmeta": float(np.abs(clf.transform(X) - g["meta"]).max()),
            "max_ddecision": float(np.abs(clf.decision_function(X) - so.decision_function(model, X)).max()),
        }
        specs = specs_for(name)
        x = P.synth_rows_device(specs, 2020, 0, n_big)
        host = P.synth.host_rows(specs, 2020, 12345, 4096)
        res["synth_bit_identical"] = bool(np.array_equal(x[:, 12345:12345 + 4096].T.cpu().numpy(), host))
        st = clf.new_status()
        o32 = clf.predict_proba_device(x, n=n_big, status=st)
        torch.cuda.synchronize()
        res["fixup_rows"] = int(st[1].item())
        res["status0"] = int(st[0].item())
        n_cmp = min(n_big, 4_000_000)
        t0 = time.time()
        o64 = clf.predict_proba_f64_device(x[:, :n_cmp + 0].contiguous() if False else x, n=n_cmp)
        torch.cuda.synchronize()
        res["f64_path_s"] = time.time() - t0
        d = (o32[:n_cmp].double() - o64).abs()
        res["max_f32_vs_f64"] = float(d.max().item())
        res["p999999_f32_vs_f64"] = float(torch.quantile(d[:1_000_000, 1], 0.999999).item())
        # oracle spot check of the f64 path on 20k rows
        idx = np.arange(0, n_cmp, n_cmp // 20000)[:20000]
        Xs = x[:, idx].T.cpu().numpy().astype(np.float64)
        res["max_f64path_vs_oracle"] = float(np.abs(o64[idx].cpu().numpy() - so.predict_proba(model, Xs)).max())
        res["max_f32path_vs_oracle"] = float(np.abs(o32[idx].double().cpu().numpy() - so.predict_proba(model, Xs)).max())
        # timing
        outbuf = torch.empty((n_big, 2), dtype=torch.float32, device="cuda")
        for _ in range(3):
            clf.predict_proba_device(x, n=n_big, out=outbuf, status=st)
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
        for a, b in evs:
            a.record()
            clf.predict_proba_device(x, n=n_big, out=outbuf, status=st)
            b.record()
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in evs)
        res["kernel_ms_median"] = ms[len(ms) // 2]
        res["rows_per_s"] = n_big / (ms[len(ms) // 2] * 1e-3)
        res["hbm_frac"] = res["rows_per_s"] * 76 / 6545.6e9
        res["n_sv"] = int(clf.info.n_support_vectors)
        out[name] = res
        print(name, json.dumps(res), flush=True)
        del x, o32, o64, outbuf
        torch.cuda.empty_cache()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "probe.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
