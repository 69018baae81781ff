"""Round-2 ncu outputs (tools/prof_round.sh -> gpurun_out/) into the tracked summaries under profiles/ (run on the CPU box).

    python tools/summarize_r2.py

Writes profiles/r2_launches.{md,csv}, profiles/r2_{mma,stream,forest}_kernel.md + *_raw.csv, trimmed SASS extracts
profiles/r2_{mma,stream,forest}_sass.txt, and the measured DRAM traffic per launch into profiles/traffic.json.
"""
import collections
import csv
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
SRC = os.path.join(ROOT, "gpurun_out")
KEYS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__warps_eligible.avg.per_cycle_active",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
] + [f"smsp__average_warps_issue_stalled_{k}_per_issue_active.ratio" for k in (
    "wait", "math_pipe_throttle", "short_scoreboard", "long_scoreboard", "barrier", "not_selected", "no_instruction",
    "dispatch_stall", "mio_throttle", "sleeping", "branch_resolving")]
SCALE = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}


def launches(csv_name, command):
    rows = [r for r in csv.reader(open(os.path.join(SRC, csv_name))) if len(r) > 5]
    ci = {h: i for i, h in enumerate(rows[0])}
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[ci["Metric Value"]].replace(",", ""))
        except ValueError:
            continue
        unit = r[ci["Metric Unit"]]
        ms = v * {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0}.get(unit, 1e-3)
        a = agg.setdefault(re.sub(r"\(.*", "", r[ci["Kernel Name"]]), [0, 0.0])
        a[0] += 1
        a[1] += ms
    tot = sum(v[1] for v in agg.values())
    lines = [f"# ncu launch list of `{command}` (round 2)", "",
             "`ncu --metrics gpu__time_duration.sum --clock-control none -c 400`, first 400 launches of the process (model "
             "creation incl. the guard probe, synthetic rows, warm-up and the timed loop of the main leg); times are "
             "cold-cache and serialised: compare shares, not absolutes.", "",
             "| kernel | launches | total ms | mean ms | share |", "|---|---:|---:|---:|---:|"]
    for k, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"| `{k[:100]}` | {c} | {ms:.3f} | {ms / c:.4f} | {100 * ms / tot:.1f} % |")
    open(os.path.join(OUT, "r2_launches.md"), "w").write("\n".join(lines) + "\n")
    subprocess.run(["cp", os.path.join(SRC, csv_name), os.path.join(OUT, "r2_launches.csv")], check=True)


def ncu_page(rep, page):
    return subprocess.run(["ncu", "-i", os.path.join(SRC, rep), "--page", page, "--csv"], capture_output=True,
                          text=True).stdout


def full(rep, tag, title, command, n_rows, algo_bytes, traffic_key, model, sass_patterns, note=""):
    raw = ncu_page(rep, "raw")
    open(os.path.join(OUT, f"r2_{tag}_raw.csv"), "w").write(raw)
    rows = list(csv.reader(raw.splitlines()))
    m = {h: (rows[2][i], rows[1][i]) for i, h in enumerate(rows[0])}
    out = [f"# ncu --set full, {title} (round 2)", "",
           f"`{command}` ({n_rows} synthetic rows, model `{model}`; raw metric dump: `r2_{tag}_raw.csv`).", "",
           "| metric | value | unit |", "|---|---:|---|"]
    out += [f"| `{k}` | {m[k][0]} | {m[k][1]} |" for k in KEYS if k in m]
    tb = sum(float(m[k][0].replace(",", "")) * SCALE[m[k][1]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    out += ["", f"DRAM traffic per launch: {tb / 1e6:.1f} MB = **{tb / n_rows:.1f} B/row** against {algo_bytes} B/row "
                f"algorithmic (68 read + 8 written).", ""]
    if note:
        out += [note, ""]
    if traffic_key:
        path = os.path.join(OUT, "traffic.json")
        t = json.load(open(path)) if os.path.exists(path) else {}
        t[traffic_key] = {"model": model, "rows": n_rows, "dram_bytes": tb,
                          "source": f"profiles/r2_{tag}_raw.csv (ncu --set full, dram__bytes_read.sum + "
                                    f"dram__bytes_write.sum, one launch)"}
        json.dump(t, open(path, "w"), indent=1)
    # per-instruction page: hottest SASS lines and the instructions that prove the hardware path
    src = ncu_page(rep, "source")
    srows = list(csv.reader(src.splitlines()))
    hdr_at = next(i for i, r in enumerate(srows) if "Source" in r and "# Samples" in r)
    ci = {h: i for i, h in enumerate(srows[hdr_at])}
    data = srows[hdr_at + 1:]
    samples = [int(r[ci["# Samples"]] or 0) for r in data]
    execd = [int(r[ci["Instructions Executed"]] or 0) for r in data]
    tot = max(sum(samples), 1)
    mix = collections.Counter()
    for r, e in zip(data, execd):
        op = r[ci["Source"]].split()
        op = [w for w in op if not w.startswith("@")]
        if op:
            mix[op[0].split(".")[0]] += e
    etot = max(sum(mix.values()), 1)
    out += ["Executed warp-instructions by opcode (top 14): " +
            ", ".join(f"`{k}` {100 * v / etot:.1f} %" for k, v in mix.most_common(14)), ""]
    open(os.path.join(OUT, f"r2_{tag}_kernel.md"), "w").write("\n".join(out) + "\n")
    keep = set()
    pat = re.compile(sass_patterns)
    for i, r in enumerate(data):
        if pat.search(r[ci["Source"]]):
            keep.update(range(max(0, i - 1), min(len(data), i + 2)))
    top = sorted(range(len(data)), key=lambda i: -samples[i])[:40]
    keep.update(top)
    lines = [f"# SASS extract of {title}: lines matching /{sass_patterns}/ (+-1) and the 40 most-sampled instructions",
             f"# from {rep} (ncu --page source); columns: index, samples (share), executed, instruction", ""]
    prev = None
    for i in sorted(keep):
        if prev is not None and i != prev + 1:
            lines.append("    ...")
        lines.append(f"{i:5d} {samples[i]:7d} ({100 * samples[i] / tot:4.1f} %) {execd[i]:10d}  {data[i][ci['Source']]}")
        prev = i
    open(os.path.join(OUT, f"r2_{tag}_sass.txt"), "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    launches("r2_bench_launches.csv", "python bench.py --steps 2 --warmup 3")
    full("r2_mma_bench.ncu-rep", "mma", "tensor-core kernel `hfe_mma_kernel<stumps>` (default engine), bench model and size",
         "ncu --set full --clock-control none --import-source on -k regex:hfe_mma -s 2 -c 1 python tools/ncu_target.py",
         10_000_000, 76, "fused", "retrained",
         r"UTCHMMA|UTCBAR|LDTM|UTCATOM|MUFU\.EX2|UBLKCP|SYNCS|FFMA2|ELECT|R2UR")
    full("r2_stream_bench.ncu-rep", "stream", "streaming pass `hfe_stream_kernel<stumps>` (stack without the SVC member)",
         "ncu --set full --clock-control none --import-source on -k regex:hfe_stream -s 2 -c 1 python tools/ncu_target.py",
         10_000_000, 76, "stream", "retrained without SVC", r"UTMALDG|SYNCS|MUFU|STG")
    full("r2_forest_bench.ncu-rep", "forest", "shared-memory forest walk `hfe_forest_kernel` on the 2000-tree stress forest",
         "ncu --set full --clock-control none --import-source on -k regex:hfe_forest -s 1 -c 1 python tools/ncu_target.py",
         2_000_000, 76, "forest", "stress forest (2000 trees, no SVC)", r"UBLKCP|SYNCS")
