"""bench.py -- patient-rows/s of the stacked ensemble's predict_proba (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--rows R] [--model retrained|published|...]
                    [--impl reference] [--legs all|none|svc_free,stress,perm]

A *step* is one pass of the hot path over one batch of synthetic patient rows already resident in HBM as a
column-major float32 matrix when the timed region starts.  ``value`` = rows all ranks scored / max-over-ranks device
time.

  N = 1   BASELINE config 2: 10 M rows on one GPU.
  N > 1   BASELINE config 3 (torchrun, one rank per GPU): 1 B rows in total, row-sharded contiguously
          (``shard.shard_range``), "strong" scaling; rows are a pure function of the global row index and there is no
          collective inside the timed region.  ``--rows R`` overrides both with R rows per GPU ("weak").

The JSON line also carries
  parity       the TIMED outputs of every rank checked against the CPU oracle on >= 4096 strided rows of its shard
               (regenerated on the host from the global row index): max |dp|, labels exact;
  checksums    CRC-32 of the output bytes of eight fixed 1 M-row windows (global rows k * 125 M): the same rows give the
               same bytes whatever the GPU count, so the N = 1 line and the N = 2/4/8 lines must agree on every window;
  e2e          the same metric through the call the reference makes: ``clf.predict_proba(X)`` with X a pageable
               row-major float64 numpy matrix (load_data_public.py:11-14), host->device and device->host copies inside
               the timed region; ``e2e_pinned_f32`` / ``e2e_u8_cohort`` are the same path fed from pinned float32 / uint8;
  roofline     algorithmic bytes (76 B/row: 17 float32 in + 2 float32 out) / CUDA-event time of the scoring kernel,
               against the measured HBM copy bandwidth in MEASURED_PEAKS.json;
  fma_roofline the bound that binds the stack with its SVC member;
  legs         the other BASELINE configurations under the same clock: "svc_free" (streaming pass, the HBM-bound part),
               "stress" (config 4: 2000 trees, depth 16) and "perm_importance" (config 5), each with its own numbers;
  cpu_baseline scikit-learn's own predict_proba (the reference's implementation of this path) on the host cores, on a
               bounded sample of the same rows, rank 0 and N = 1 only.

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
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
BYTES_PER_ROW = 17 * 4 + 2 * 4          # algorithmic bytes per patient row
METRIC = "patient-rows/sec ensemble predict_proba"
SEED = 2020
ROWS_CONFIG2 = 10_000_000
ROWS_CONFIG3 = 1_000_000_000
WINDOW_STRIDE = 125_000_000             # checksum windows start at global rows k * 125 M (the 8-GPU shard starts)
WINDOW_ROWS = 1_000_000
N_WINDOWS = 8


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
    X = pkg.synth.host_rows(specs, SEED, 0, sample_rows).astype(np.float64)
    chunks = np.array_split(X, max(threads * 4, 1))
    stack.predict_proba(X[:256])
    best = None
    for _ in range(passes):
        t0 = time.perf_counter()
        Parallel(n_jobs=threads, prefer="threads")(delayed(stack.predict_proba)(c) for c in chunks)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return sample_rows / best, best


def workload_name(args, arrays, world, rows_total):
    if args.rows:
        cfg = f"{rows_total} synthetic patient rows ({args.rows} per GPU)"
    elif world == 1:
        cfg = "BASELINE config 2: 10M synthetic patient rows on 1xB200"
    else:
        cfg = f"BASELINE config 3: 1B synthetic patient rows row-sharded across {world}xB200"
    return f"{cfg}; {args.model} stack ({describe(arrays)}); predict_proba"


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
    world = args.gpus
    rows_total = args.rows * world if args.rows else (ROWS_CONFIG2 if world == 1 else ROWS_CONFIG3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "rows/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak" if (args.rows or world == 1) else "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args, arrays, world, rows_total),
                   "model": args.model, "rows_per_step": sample,
                   "note": "bounded sample of the same workload on the host cores (scikit-learn, the reference path)"},
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": cores, "kind": "reference",
                         "sample": f"{sample} synthetic rows per step, joblib threads over {cores} cores, "
                                   "scikit-learn 1.9 StackingClassifier.predict_proba rebuilt from the same arrays"},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- pieces of the GPU arm ---------------------------------------------------------------------------------------------

class Ctx:
    """What every leg needs: package, torch, ranks, the device and reductions over ranks."""

    def __init__(self, pkg, torch, rank, local_rank, world):
        self.pkg, self.torch, self.rank, self.local_rank, self.world = pkg, torch, rank, local_rank, world
        self.dev = f"cuda:{local_rank}"

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.torch.distributed.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, value, op="max"):
        if self.world == 1:
            return float(value)
        dist = self.torch.distributed
        t = self.torch.tensor([float(value)], dtype=self.torch.float64, device=self.dev)
        dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN, "sum": dist.ReduceOp.SUM}[op])
        return float(t.item())


def timed_device_loop(ctx, clf, x, n, out, status, steps, warmup):
    """K scoring calls on the current stream bracketed by CUDA events; returns (total ms max over ranks, per-call ms)."""
    torch = ctx.torch
    for _ in range(warmup):
        clf.predict_proba_device(x, n=n, out=out, status=status)
    ctx.barrier()
    # two events around the K calls: an event recorded between calls costs ~9 us of device idle time per step on the
    # 150 us streaming pass (measured), so the per-call figure is total / K
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        clf.predict_proba_device(x, n=n, out=out, status=status)
    e1.record()
    ctx.barrier()
    local_ms = e0.elapsed_time(e1)
    total_ms = ctx.reduce(local_ms, "max")
    return total_ms, [local_ms / steps] * steps


def parity_block(ctx, clf, arrays, specs, x, out, begin, n, sample=4096):
    """The timed outputs against the CPU oracle on strided rows of this rank's shard (rows regenerated on the host)."""
    from oracle import stack_oracle as so            # the checker; never inside a timed region

    torch = ctx.torch
    step = max(n // sample, 1)
    local = np.arange(0, n, step)[:sample]
    X = ctx.pkg.synth.host_rows(specs, SEED, 0, 0, rows=begin + local).astype(np.float64)
    sel = torch.from_numpy(local).to(ctx.dev)
    same_rows = bool(np.array_equal(x[:, sel].T.cpu().numpy().astype(np.float64), X))   # host mirror == device rows
    z = so.decision_function(arrays, X)
    p1 = 1.0 / (1.0 + np.exp(-z))
    got = out[sel].cpu().numpy().astype(np.float64)
    err = float(np.abs(got - np.stack([1.0 - p1, p1], axis=1)).max())
    xs = torch.zeros((x.shape[0], (len(local) + 3) & ~3), dtype=torch.float32, device=ctx.dev)
    xs[:, :len(local)] = x[:, sel]
    lab = torch.empty(len(local), dtype=torch.uint8, device=ctx.dev)
    clf.score_device(xs, n=len(local), labels=lab)
    torch.cuda.synchronize()
    labels_ok = bool(np.array_equal(lab.cpu().numpy(), (z > 0).astype(np.uint8)))
    return {"max_abs_dp": ctx.reduce(err, "max"), "labels_exact": bool(ctx.reduce(float(labels_ok and same_rows), "min")),
            "rows": int(ctx.reduce(len(local), "sum")), "tolerance": 1e-6,
            "against": "oracle/stack_oracle.py (numpy float64 restatement, pinned to scikit-learn) on the timed outputs"}


def checksum_windows(ctx, clf, specs, out, begin, end, rows_total):
    """CRC-32 of the float32 output bytes of fixed global row windows.  A window that lies inside this rank's shard is
    read from the timed output buffer; windows no shard holds (N = 1 scores only the first 10 M rows) are generated and
    scored on their own by rank k mod N."""
    torch = ctx.torch
    sums = torch.zeros(N_WINDOWS, dtype=torch.int64, device=ctx.dev)
    for k in range(N_WINDOWS):
        w0, w1 = k * WINDOW_STRIDE, k * WINDOW_STRIDE + WINDOW_ROWS
        inside_total = w1 <= rows_total
        if inside_total and begin <= w0 and w1 <= end:
            piece = out[w0 - begin:w1 - begin]
        elif not inside_total and k % ctx.world == ctx.rank:
            xw = ctx.pkg.synth_rows_device(specs, SEED, w0, WINDOW_ROWS, device=ctx.local_rank)
            piece = clf.predict_proba_device(xw, n=WINDOW_ROWS)
            torch.cuda.synchronize()
        else:
            continue                                           # another rank owns it (or it straddles a shard boundary)
        sums[k] = zlib.crc32(piece.contiguous().cpu().numpy().tobytes())
    if ctx.world > 1:
        torch.distributed.all_reduce(sums)
    return {"crc32": [f"{int(v):08x}" for v in sums.cpu().tolist()], "window_rows": WINDOW_ROWS,
            "window_starts": [k * WINDOW_STRIDE for k in range(N_WINDOWS)],
            "note": "same global rows -> same bytes for any GPU count: compare with the N = 1 line"}


def e2e_legs(ctx, clf, x, out, n, args):
    """End to end through the public host API, host<->device copies inside the timed region."""
    torch = ctx.torch
    rows = min(n, int(args.e2e_rows))
    steps = max(2, min(args.steps, 5))
    res = {}
    # (1) the reference's call: clf.predict_proba(X), X pageable row-major float64 (load_data_public.py:11-14)
    X64 = x[:, :rows].T.contiguous().cpu().numpy().astype(np.float64)
    for _ in range(2):
        proba = clf.predict_proba(X64)
    ctx.barrier()
    calls = []
    t0 = time.perf_counter()
    for _ in range(steps):
        c0 = time.perf_counter()
        proba = clf.predict_proba(X64)
        calls.append(time.perf_counter() - c0)
    dt = ctx.reduce((time.perf_counter() - t0) / steps, "max")
    diff = float(np.abs(proba[:, 1] - out[:rows, 1].cpu().numpy().astype(np.float64)).max())
    # the library sends a chunk as bytes when all its values are integers in [0, 255] (lossless), else as float32
    as_bytes = bool(((X64 == np.rint(X64)) & (X64 >= 0) & (X64 <= 255)).all())
    res["e2e"] = {"value": rows * ctx.world / dt, "unit": "rows/s", "h2d_bytes_per_step": rows * 17 * (1 if as_bytes else 4),
                  "wire_format": "uint8 (every value an integer in [0, 255]: lossless)" if as_bytes else "float32",
                  "d2h_bytes_per_step": rows * 2 * 4, "rows_per_step": rows, "steps": steps,
                  "host_input_bytes_per_step": rows * 17 * 8, "host_output_bytes_per_step": rows * 2 * 8,
                  "api": "HFStackB200.predict_proba(X): X numpy float64 [rows,17] C-order in pageable memory -> numpy "
                         "float64 [rows,2]; hfe_predict_host narrows on host threads into pinned chunks (bytes when the "
                         "chunk is an integer cohort, else float32: what crosses PCIe), 3 chunks in flight over 3 streams, "
                         "ingest + scoring + D2H per chunk",
                  "max_abs_diff_vs_device_path": diff, "ms_per_call_rank0": [round(c * 1e3, 2) for c in calls]}
    del X64, proba
    # (2) the same device path fed from pinned float32 / uint8 tensors (no host-side narrowing)
    x_host = torch.empty((rows, 17), dtype=torch.float32).pin_memory()
    x_host.copy_(x[:, :rows].T)
    out_host = torch.empty((rows, 2), dtype=torch.float32).pin_memory()
    for key, xh, bpr in (("e2e_pinned_f32", x_host, 68), ("e2e_u8_cohort", None, 17)):
        if xh is None:
            xh = torch.empty((rows, 17), dtype=torch.uint8).pin_memory()
            xh.copy_(x[:, :rows].T)
            lossless = bool((xh.float() == x_host).all().item())
        for _ in range(2):
            clf.predict_proba_pinned(xh, out_host)
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            clf.predict_proba_pinned(xh, out_host)
        torch.cuda.synchronize()
        dt = ctx.reduce((time.perf_counter() - t0) / steps, "max")
        res[key] = {"value": rows * ctx.world / dt, "unit": "rows/s", "h2d_bytes_per_step": rows * bpr,
                    "d2h_bytes_per_step": rows * 8, "rows_per_step": rows, "steps": steps,
                    "h2d_gb_per_s_per_gpu": rows * bpr / dt / 1e9,
                    "api": "HFStackB200.predict_proba_pinned (pinned rows in, pinned probabilities out)",
                    "max_abs_diff_vs_device_path": float((out_host[:4096].to(ctx.dev) - out[:4096]).abs().max().item())}
    # the host link by itself: every rank copies the same pinned float32 rows host -> device at the same time, nothing else
    # running -- the ceiling any end-to-end figure at this GPU count is to be read against
    dst = torch.empty_like(x_host, device=ctx.dev)
    dst.copy_(x_host, non_blocking=True)
    torch.cuda.synchronize()
    ctx.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        dst.copy_(x_host, non_blocking=True)
    torch.cuda.synchronize()
    dt = ctx.reduce((time.perf_counter() - t0) / steps, "max")
    res["host_link"] = {"plain_h2d_gb_per_s_per_gpu": x_host.numel() * 4 / dt / 1e9,
                        "plain_h2d_gb_per_s_all_gpus": x_host.numel() * 4 * ctx.world / dt / 1e9,
                        "host_cpus": len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else None,
                        "note": "concurrent cudaMemcpyAsync of 680 MB pinned float32 per rank; host_cpus = CPUs this rank is bound to"}
    del dst
    res["e2e_u8_cohort"]["lossless"] = lossless
    res["e2e_u8_cohort"]["note"] = ("the cohort held as uint8 (every schema variable is a small non-negative integer); "
                                    "not the headline: the reference takes float arrays")
    return res


def leg_svc_free(ctx, arrays, x, n, hbm_peak):
    """The stack without its SVC member -- the part of the ensemble that is HBM-bound -- through the streaming pass."""
    torch = ctx.torch
    clf = ctx.pkg.HFStackB200.from_arrays(ctx.pkg.synth_models.without_member(arrays, "svc"), device=ctx.local_rank)
    rows = min(n, ROWS_CONFIG2)
    out = torch.empty((rows, 2), dtype=torch.float32, device=ctx.dev)
    st = clf.new_status()
    k = 40                                    # 0.15 ms per pass: enough passes that the first ones do not weigh
    total_ms, per = timed_device_loop(ctx, clf, x, rows, out, st, steps=k, warmup=5)
    ms = total_ms / k
    achieved = BYTES_PER_ROW * rows / (sum(per) / len(per) * 1e-3) / 1e9
    clf.close()
    return {"workload": "same stack without the SVC member (boosted stumps + logistic -> combiner), streaming pass",
            "value": rows * ctx.world / (ms * 1e-3), "unit": "rows/s", "rows_per_gpu": rows, "steps": k, "warmup": 5,
            "ms_per_step": ms, "gpu_launches": 2 * k * ctx.world,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved / hbm_peak, "kernel": "hfe_stream_kernel + fix-up launch per call",
                         "bytes_per_row": BYTES_PER_ROW}}


def leg_stress(ctx, arrays, x, n, hbm_peak):
    """BASELINE config 4: 2000 trees, depth <= 16, node tables far beyond shared memory."""
    torch = ctx.torch
    stress = ctx.pkg.synth_models.stress_stack(arrays, with_svc=False)
    clf = ctx.pkg.HFStackB200.from_arrays(stress, device=ctx.local_rank)
    rows = min(n, 2_000_000)                      # 136 MB of features per pass: larger than the 126 MB L2
    out = torch.empty((rows, 2), dtype=torch.float32, device=ctx.dev)
    st = clf.new_status()
    total_ms, per = timed_device_loop(ctx, clf, x, rows, out, st, steps=3, warmup=3)
    ms = total_ms / 3
    achieved = BYTES_PER_ROW * rows / (sum(per) / len(per) * 1e-3) / 1e9
    info = clf.info
    res = {"workload": f"BASELINE config 4: stress forest, {info.n_trees} trees depth <= {info.max_tree_depth}, "
                       f"{info.n_tree_nodes} nodes, + logistic -> combiner",
           "value": rows * ctx.world / (ms * 1e-3), "unit": "rows/s", "rows_per_gpu": rows, "steps": 3, "warmup": 3,
           "ms_per_step": ms, "tree_walks_per_s": rows * ctx.world * int(info.n_trees) / (ms * 1e-3),
           "table_bytes": int(info.device_bytes), "gpu_launches": 6 * ctx.world,
           "f64_reevaluated_rows_last_step": int(st[1].item()),
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                        "frac": achieved / hbm_peak, "bytes_per_row": BYTES_PER_ROW,
                        "note": "32 000 dependent node visits per row out of tables that live in L2/L1: not an "
                                "HBM-bound configuration; see profiles/ for the L1/L2 counters"}}
    clf.close()
    return res


def leg_perm(ctx, clf, specs, n):
    """BASELINE config 5: permutation importance, 17 columns x 5 repeats + baseline = 86 passes over 1 M rows.
    Every rank holds the same cohort (global rows [0, 1 M)) and scores its share of the 86 passes."""
    torch = ctx.torch
    pkg = ctx.pkg
    rows = min(n, 1_000_000)
    x = pkg.synth_rows_device(specs, SEED, 0, rows, device=ctx.local_rank)
    lab = torch.empty(rows, dtype=torch.uint8, device=ctx.dev)
    clf.score_device(x, n=rows, labels=lab)
    flip = torch.from_numpy(pkg.synth.hash_u32(99, np.arange(rows), 251) < 2 ** 29).to(ctx.dev)   # 12.5 % label noise
    y = (lab ^ flip.to(torch.uint8)).cpu().numpy().astype(np.float64)
    res = {"workload": "BASELINE config 5: permutation importance, 86 predict passes over 1M rows", "rows": rows,
           "passes": 86, "n_gpus": ctx.world}
    for mode in ("sklearn", "device"):
        pkg.inspection.permutation_importance(clf, x, y, n_repeats=1, random_state=0, rank=ctx.rank, world=ctx.world,
                                              shuffle=mode)
        ctx.barrier()
        t0 = time.perf_counter()
        r = pkg.inspection.permutation_importance(clf, x, y, n_repeats=5, random_state=SEED, rank=ctx.rank,
                                                  world=ctx.world, shuffle=mode)
        torch.cuda.synchronize()
        dt = ctx.reduce(time.perf_counter() - t0, "max")
        res[f"seconds_{mode}_shuffles"] = dt
        res[f"rows_scored_per_s_{mode}_shuffles"] = 86 * rows / dt
        if mode == "sklearn":
            res["baseline_accuracy"] = float(r.baseline_score)
            top = np.argsort(-r.importances_mean)[:3]
            res["top_features"] = [[int(i), round(float(r.importances_mean[i]), 6)] for i in top]
    return res


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
    ctx = Ctx(pkg, torch, rank, local_rank, world)
    clf = pkg.HFStackB200.from_arrays(arrays, device=local_rank)
    if int(clf.info.n_support_vectors):
        clf.set_rbf_engine(args.engine)
    if args.rows:
        n_total, scaling = int(args.rows) * world, "weak"
    else:
        n_total, scaling = (ROWS_CONFIG2, "weak") if world == 1 else (ROWS_CONFIG3, "strong")
    begin, end = shard.shard_range(n_total, rank, world)
    n = end - begin
    x = pkg.synth_rows_device(specs, SEED, begin, n, device=local_rank)
    out = torch.empty((n, 2), dtype=torch.float32, device=ctx.dev)
    status = clf.new_status()

    # ---- the timed region: K scoring calls, inputs resident in HBM -----------------------------------------------------
    for _ in range(args.warmup):
        clf.predict_proba_device(x, n=n, out=out, status=status)
    ctx.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ctx.barrier()
    t_wall0 = time.time()
    evs[0].record()
    for i in range(args.steps):
        clf.predict_proba_device(x, n=n, out=out, status=status)
        evs[i + 1].record()
    ctx.barrier()
    t_wall1 = time.time()
    total_ms = evs[0].elapsed_time(evs[-1])
    per_launch = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    total_ms = ctx.reduce(total_ms, "max")
    clf.check(status)                                    # NaN/Inf or a work-list overflow on synthetic rows: fail loudly
    fixups = int(status[1].item())
    ms_per_step = total_ms / args.steps
    value = n_total / (ms_per_step * 1e-3)

    parity = parity_block(ctx, clf, arrays, specs, x, out, begin, n)
    checksums = checksum_windows(ctx, clf, specs, out, begin, end, n_total)
    e2e = e2e_legs(ctx, clf, x, out, n, args)
    hbm_peak, peak_src, sm_max = measured_peaks()
    legs = {}
    want = set(("svc_free", "stress", "perm") if args.legs == "all" else
               () if args.legs == "none" else args.legs.split(","))
    n_sv = int(clf.info.n_support_vectors)
    if "svc_free" in want and n_sv:
        legs["svc_free"] = leg_svc_free(ctx, arrays, x, n, hbm_peak)
    if "stress" in want:
        legs["stress"] = leg_stress(ctx, arrays, x, n, hbm_peak)
    if "perm" in want:
        legs["perm_importance"] = leg_perm(ctx, clf, specs, n)

    if rank != 0:
        return
    avg_launch_ms = sum(per_launch) / len(per_launch)
    achieved = BYTES_PER_ROW * n / (avg_launch_ms * 1e-3) / 1e9
    # lane-FMAs the RBF loop really executes per row (the prefix-shared table layout shares partial dot products, so
    # this is below the flat n_sv * 18); the library reports it from the table it built
    fma_per_row = int(clf.info.rbf_fma_per_row)
    fma_per_row_flat = (n_sv + 3) // 4 * 4 * 18
    sm_mhz = (clocks or {}).get("sm_mhz") or sm_max
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    fma_peak = sms * 128 * sm_mhz * 1e6
    traffic, traffic_profile = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            t = json.load(fh)
        cap = t.get("fused" if n_sv else "stream")
        if cap and int(clf.info.tree_mode) != 2:
            # one `ncu --set full` capture of this kernel; it is this run's traffic only when model and row count match
            if cap.get("model") == args.model and int(cap.get("rows", 0)) == n:
                traffic = cap["dram_bytes"]
            traffic_profile = {"dram_bytes_per_row": cap["dram_bytes"] / cap["rows"], "model": cap.get("model"),
                               "rows": cap.get("rows"), "source": cap.get("source")}
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, arrays, world, n_total),
                   "model": args.model, "rows_per_gpu": n, "rows_total": n_total, "layout": "column-major f32 [17][n]",
                   "l2": f"inputs larger than L2 ({n * 68 / 1e6:.0f} MB per pass vs 126 MB)",
                   "f64_reevaluated_rows_last_step": fixups},
        "clocks": clocks,
        "parity": parity,
        "checksums": checksums,
        "e2e": e2e["e2e"],
        "e2e_pinned_f32": e2e["e2e_pinned_f32"],
        "e2e_u8_cohort": e2e["e2e_u8_cohort"],
        "host_link": e2e["host_link"],
        "gpu_launches": 2 * args.steps * world,      # scoring pass + work-list fix-up per step and rank
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "traffic_from_profile": traffic_profile,
                     "peak_source": peak_src,
                     "kernel": ("hfe_score_kernel (fused pass" if n_sv or int(clf.info.tree_mode) == 2 else
                                "hfe_stream_kernel (streaming pass") + "; the work-list fix-up launch that follows takes "
                               "microseconds)",
                     "bytes_per_row": BYTES_PER_ROW, "avg_launch_ms": avg_launch_ms,
                     "note": ("the RBF-SVC member makes this pass FP32-FMA / MUFU bound, not HBM bound: see fma_roofline"
                              if n_sv else "no SVC member: this configuration is the HBM-bound part of the ensemble")},
    }
    engine = int(clf.info.rbf_engine)
    line["config"]["rbf_engine"] = {0: "none (no SVC member)", 1: "simt (FFMA2 loop)", 2: "tensor cores (tcgen05.mma + TMEM)"}[engine]
    if engine == 2:
        # the tensor-core engine leaves one exponential per (row, support-vector column) on the MUFU pipe, and as many
        # float32 values to read back from TMEM: both paths move 16 per clock and SM
        exps = int(clf.info.rbf_exp_per_row)
        peak = sms * 16 * sm_mhz * 1e6
        line["roofline"]["kernel"] = "hfe_mma_kernel (tcgen05 RBF contraction + MUFU epilogue; the fix-up launch that follows takes microseconds)"
        line["roofline"]["note"] = "the RBF-SVC member makes this pass MUFU / TMEM-read bound, not HBM bound: see mufu_roofline"
        line["mufu_roofline"] = {"bound": "MUFU.EX2 pipe = TMEM read path (16 values per clock and SM)",
                                 "achieved": exps * n / (avg_launch_ms * 1e-3) / 1e12, "peak": peak / 1e12,
                                 "unit": "T exp/s", "frac": exps * n / (avg_launch_ms * 1e-3) / peak,
                                 "exp_per_row": exps, "sm_mhz_used": sm_mhz, "sms": sms}
    if n_sv and engine == 1:
        line["fma_roofline"] = {"bound": "fp32 fma pipe", "achieved": fma_per_row * n / (avg_launch_ms * 1e-3) / 1e12,
                                "peak": fma_peak / 1e12, "unit": "T lane-FMA/s",
                                "frac": fma_per_row * n / (avg_launch_ms * 1e-3) / fma_peak,
                                "fma_per_row": fma_per_row, "fma_per_row_flat_layout": fma_per_row_flat,
                                "frac_if_counted_as_flat_layout": fma_per_row_flat * n / (avg_launch_ms * 1e-3) / fma_peak,
                                "sm_mhz_used": sm_mhz, "sms": sms,
                                "note": "achieved counts the lane-FMAs executed; the prefix-shared table removes a third "
                                        "of the flat layout's FMAs, which leaves the FMA and MUFU pipes co-critical"}
    if legs:
        line["legs"] = legs
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
    ap.add_argument("--rows", type=int, default=0,
                    help="rows per GPU per step (default: 10 M at N = 1, 1 B in total at N > 1)")
    ap.add_argument("--model", default="retrained",
                    choices=["retrained", "published", "published_nosvc", "stress", "stress_svc"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-rows", type=int, default=10_000_000)
    ap.add_argument("--cpu-rows", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--legs", default="all", help="all | none | comma list of svc_free,stress,perm")
    ap.add_argument("--engine", default="auto", choices=["auto", "simt", "tensor"],
                    help="what evaluates the RBF sum: tensor cores (tcgen05) or the SIMT FFMA2 loop; auto = tensor when it fits")
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
