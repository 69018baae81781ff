"""bench.py -- patient-rows/s of the stacked ensemble's predict_proba (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--rows R] [--model retrained|published] [--impl reference]

A *step* is one pass of the hot path over one batch: ``--rows`` synthetic patient rows per GPU (default 10 M = config
"same ensemble, 10M synthetic patient rows, predict_proba on 1xB200"), already resident in HBM as a column-major
float32 matrix when the timed region starts.  ``value`` = rows all ranks scored / max-over-ranks device time.
With N > 1 (torchrun, one rank per GPU) every rank scores its own shard of N*R rows ("weak" scaling; rows are a pure
function of the global row index); there is no collective inside the timed region.

The JSON line also carries
  e2e          the same metric through the public host API (``HFStackB200.predict_proba_pinned``): pinned host float32
               rows in, pinned host probabilities out, H2D + ingest + kernel + D2H inside the timed region;
  roofline     algorithmic bytes (76 B/row: 17 float32 in + 2 float32 out) / CUDA-event time of the fused kernel,
               against the measured HBM copy bandwidth in MEASURED_PEAKS.json;
  fma_roofline the bound that actually binds the full ensemble: float32 lane-FMAs issued per
               second against 148 SMs x 128 lanes x the SM clock observed during the run;
  cpu_baseline scikit-learn's own predict_proba (the reference's implementation of this path) on the host cores, on a
               bounded sample of the same rows, rank 0 and N=1 only.

``--impl reference`` times only that scikit-learn path (all host threads) and prints the same line shape.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
BYTES_PER_ROW = 17 * 4 + 2 * 4          # algorithmic bytes per patient row
METRIC = "patient-rows/sec ensemble predict_proba"


def load_model(name):
    """(package, ModelArrays, column specs of the synthetic rows) for a --model choice."""
    pkg = importlib.import_module("machine-learning-replications_b200")
    base = "retrained" if name == "retrained" else "published"
    arrays = pkg.export.read_blob(open(os.path.join(GOLDEN, f"{base}_model.blob"), "rb").read())
    if base == "published":
        specs = pkg.synth.schema_specs()
    else:
        cand = pkg.synth.candidate_specs()
        specs = [cand[c] for c in np.load(os.path.join(GOLDEN, "retrained_columns.npy"))]
    if name == "published_nosvc":          # the members that are genuinely HBM-bound
        arrays = pkg.synth_models.without_member(arrays, "svc")
    elif name == "stress":                 # BASELINE.json config 4: 2000 trees, depth 16, node tables >> shared memory
        arrays = pkg.synth_models.stress_stack(arrays, with_svc=False)
    elif name == "stress_svc":
        arrays = pkg.synth_models.stress_stack(arrays, with_svc=True)
    return pkg, arrays, specs


def describe(arrays):
    parts = []
    for m in arrays.members:
        if m.kind == "svc":
            parts.append(f"scaler+RBF-SVC {m.sv.shape[0]} SVs")
        elif m.kind == "gbt":
            nodes = sum(t.left.shape[0] for t in m.trees)
            parts.append(f"{len(m.trees)} boosted trees ({nodes} nodes)")
        else:
            parts.append("L1 logistic")
    return " + ".join(parts) + " -> logistic combiner"


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peaks = json.load(fh)
        return float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(peaks.get("sm_max_mhz", 1965.0))
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        inside = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows[-3:]]
        sm = sorted(float(r[1]) for r in inside if r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in inside if r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in inside for n, v in zip(names, r[4:8]) if v.lower().startswith("active")})
        pw = [float(r[3]) for r in inside if r[3].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(inside), "power_w_max": max(pw) if pw else None}


def sklearn_rows_per_s(arrays, specs, pkg, sample_rows, threads, passes=1):
    """scikit-learn's StackingClassifier.predict_proba (rebuilt from the same arrays) on ``threads`` host threads."""
    import warnings

    from joblib import Parallel, delayed

    from oracle import sk_rebuild

    warnings.simplefilter("ignore")
    stack = sk_rebuild.rebuild_stack(arrays)
    X = pkg.synth.host_rows(specs, 2020, 0, sample_rows).astype(np.float64)
    chunks = np.array_split(X, max(threads * 4, 1))
    stack.predict_proba(X[:256])
    best = None
    for _ in range(passes):
        t0 = time.perf_counter()
        Parallel(n_jobs=threads, prefer="threads")(delayed(stack.predict_proba)(c) for c in chunks)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return sample_rows / best, best


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    pkg, arrays, specs = load_model(args.model)
    cores = len(os.sched_getaffinity(0))
    sample = int(args.cpu_rows or min(max(20000 * cores, 20000), 2_000_000))
    times = []
    for i in range(args.warmup + args.steps):
        rps, dt = sklearn_rows_per_s(arrays, specs, pkg, sample, cores)
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    value = sample / (ms * 1e-3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "rows/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.model} stack ({describe(arrays)}), synthetic patient rows, predict_proba",
                   "model": args.model, "rows_per_step": sample,
                   "note": "bounded sample of the same workload on the host cores (scikit-learn, the reference path)"},
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": cores, "kind": "reference",
                         "sample": f"{sample} synthetic rows per step, joblib threads over {cores} cores, "
                                   "scikit-learn 1.9 StackingClassifier.predict_proba rebuilt from the same arrays"},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch

    pkg, arrays, specs = load_model(args.model)
    shard = pkg.shard
    rank, local_rank, world = shard.init_process_group()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the scoring path")
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    clf = pkg.HFStackB200.from_arrays(arrays, device=local_rank)
    rows = int(args.rows)
    n_total = rows * world
    begin, end = shard.shard_range(n_total, rank, world)
    n = end - begin
    x = pkg.synth_rows_device(specs, 2020, begin, n, device=local_rank)
    out = torch.empty((n, 2), dtype=torch.float32, device=dev)
    status = clf.new_status()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        clf.predict_proba_device(x, n=n, out=out, status=status)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    barrier()
    t_wall0 = time.time()
    evs[0].record()
    for i in range(args.steps):
        clf.predict_proba_device(x, n=n, out=out, status=status)
        evs[i + 1].record()
    barrier()
    t_wall1 = time.time()
    total_ms = evs[0].elapsed_time(evs[-1])
    per_launch = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    total_ms = shard.max_over_ranks(total_ms, world, device=dev)
    fixups = int(status[1].item())
    if int(status[0].item()) != 0:
        raise SystemExit("scoring reported non-finite input on synthetic rows")
    ms_per_step = total_ms / args.steps
    value = n_total / (ms_per_step * 1e-3)

    # ---- end to end through the public host API: pinned float32 rows in, pinned probabilities out -------------------
    e2e_rows = min(n, int(args.e2e_rows))
    x_host = torch.empty((e2e_rows, 17), dtype=torch.float32).pin_memory()
    x_host.copy_(x[:, :e2e_rows].T)
    out_host = torch.empty((e2e_rows, 2), dtype=torch.float32).pin_memory()
    for _ in range(2):
        clf.predict_proba_pinned(x_host, out_host)
    barrier()
    e2e_steps = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        clf.predict_proba_pinned(x_host, out_host)
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    e2e_s = shard.max_over_ranks(e2e_s, world, device=dev)
    e2e_value = e2e_rows * world / e2e_s
    e2e_check = float((out_host[:4096].to(dev) - out[:4096]).abs().max().item())
    # the same cohort kept as unsigned bytes (every schema variable is a small non-negative integer): 17 B/row over PCIe
    x_u8 = torch.empty((e2e_rows, 17), dtype=torch.uint8).pin_memory()
    x_u8.copy_(x[:, :e2e_rows].T)
    u8_exact = bool((x_u8.float() == x_host).all().item())
    for _ in range(2):
        clf.predict_proba_pinned(x_u8, out_host)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        clf.predict_proba_pinned(x_u8, out_host)
    torch.cuda.synchronize()
    u8_s = shard.max_over_ranks((time.perf_counter() - t0) / e2e_steps, world, device=dev)
    u8_value = e2e_rows * world / u8_s
    u8_check = float((out_host[:4096].to(dev) - out[:4096]).abs().max().item())

    if rank != 0:
        return
    hbm_peak, peak_src, sm_max = measured_peaks()
    avg_launch_ms = sum(per_launch) / len(per_launch)
    achieved = BYTES_PER_ROW * n / (avg_launch_ms * 1e-3) / 1e9
    n_sv = int(clf.info.n_support_vectors)
    # lane-FMAs the RBF loop really executes per row (the prefix-shared table layout shares partial dot products, so
    # this is below the flat n_sv * 18); the library reports it from the table it built
    fma_per_row = int(clf.info.rbf_fma_per_row)
    fma_per_row_flat = (n_sv + 3) // 4 * 4 * 18
    sm_mhz = (clocks or {}).get("sm_mhz") or sm_max
    fma_peak = 148 * 128 * sm_mhz * 1e6
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            t = json.load(fh)
        # per-kernel captures: the fused pass (stacks with the SVC member) and the streaming pass (without)
        key = "dram_bytes_per_row" if n_sv else "dram_bytes_per_row_stream"
        traffic = t[key] * n if key in t and int(clf.info.tree_mode) != 2 else None
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.model} stack ({describe(arrays)}), synthetic patient rows, predict_proba",
                   "model": args.model, "rows_per_gpu": rows, "rows_total": n_total, "layout": "column-major f32 [17][n]",
                   "l2": "inputs larger than L2 (680 MB per pass vs 126 MB)", "f64_reevaluated_rows_last_step": fixups},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "rows/s", "h2d_bytes_per_step": e2e_rows * 17 * 4,
                "d2h_bytes_per_step": e2e_rows * 2 * 4, "rows_per_step": e2e_rows, "steps": e2e_steps,
                "api": "HFStackB200.predict_proba_pinned (H2D + ingest + fused kernel + D2H, 2 streams)",
                "max_abs_diff_vs_device_path": e2e_check},
        "e2e_u8_cohort": {"value": u8_value, "unit": "rows/s", "h2d_bytes_per_step": e2e_rows * 17,
                          "d2h_bytes_per_step": e2e_rows * 2 * 4, "rows_per_step": e2e_rows, "steps": e2e_steps,
                          "lossless": u8_exact, "max_abs_diff_vs_device_path": u8_check,
                          "note": "same call with the cohort held as uint8 (not the headline e2e: the reference takes "
                                  "float arrays); shows that e2e above is bound by the 68 B/row host-to-device copy"},
        "gpu_launches": 2 * args.steps * world,      # fused pass + work-list fix-up per step and rank
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                     "kernel": ("hfe_score_kernel (fused pass" if n_sv or int(clf.info.tree_mode) == 2 else
                                "hfe_stream_kernel (streaming pass") + "; the work-list fix-up launch that follows takes "
                               "microseconds)",
                     "bytes_per_row": BYTES_PER_ROW, "avg_launch_ms": avg_launch_ms,
                     "note": ("the RBF-SVC member makes this pass FP32-FMA bound, not HBM bound: see fma_roofline"
                              if n_sv else "no SVC member: this configuration is the HBM-bound part of the ensemble")},
    }
    if n_sv:
        line["fma_roofline"] = {"bound": "fp32 fma pipe", "achieved": fma_per_row * n / (avg_launch_ms * 1e-3) / 1e12,
                                "peak": fma_peak / 1e12, "unit": "T lane-FMA/s",
                                "frac": fma_per_row * n / (avg_launch_ms * 1e-3) / fma_peak,
                                "fma_per_row": fma_per_row, "fma_per_row_flat_layout": fma_per_row_flat,
                                "frac_if_counted_as_flat_layout": fma_per_row_flat * n / (avg_launch_ms * 1e-3) / fma_peak,
                                "sm_mhz_used": sm_mhz,
                                "note": "achieved counts the lane-FMAs executed; the prefix-shared table removes a third "
                                        "of the flat layout's FMAs, which leaves the FMA and MUFU pipes co-critical"}
    if world == 1 and not args.no_cpu:
        cores = len(os.sched_getaffinity(0))
        sample = int(args.cpu_rows or min(max(20000 * cores, 20000), 2_000_000))
        rps, dt = sklearn_rows_per_s(arrays, specs, pkg, sample, cores)
        line["cpu_baseline"] = {"value": rps, "unit": "rows/s", "cores": cores, "kind": "reference",
                                "sample": f"{sample} of the same synthetic rows in {dt:.1f} s, scikit-learn 1.9 "
                                          f"predict_proba chunked over {cores} joblib threads"}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--rows", type=int, default=10_000_000, help="rows per GPU per step")
    ap.add_argument("--model", default="retrained",
                    choices=["retrained", "published", "published_nosvc", "stress", "stress_svc"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-rows", type=int, default=10_000_000)
    ap.add_argument("--cpu-rows", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:
            pass


if __name__ == "__main__":
    main()
