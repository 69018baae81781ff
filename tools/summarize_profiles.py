"""Turn the ncu outputs a gpurun call left in gpurun_out/ into the tracked summaries under profiles/ (run on the CPU box).

    python tools/summarize_profiles.py        # reads gpurun_out/launches_r1.csv, prof_r1_fused.ncu-rep, prof_r1_nosvc.ncu-rep
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
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "smsp__warps_eligible.avg.per_cycle_active", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
] + [f"smsp__average_warps_issue_stalled_{k}_per_issue_active.ratio" for k in (
    "wait", "math_pipe_throttle", "short_scoreboard", "long_scoreboard", "barrier", "not_selected", "no_instruction",
    "dispatch_stall", "mio_throttle")]
SCALE = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}


def launches():
    rows = [r for r in csv.reader(open(os.path.join(SRC, "launches_r1.csv"))) if len(r) > 5]
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
    lines = ["# ncu launch list of `python bench.py --steps 10 --warmup 3 --no-cpu` (round 1)", "",
             "`ncu --metrics gpu__time_duration.sum --clock-control none -c 400`; times are cold-cache and serialised: "
             "compare shares.", "", "| kernel | launches | total ms | share |", "|---|---:|---:|---:|"]
    for k, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"| `{k[:90]}` | {c} | {ms:.3f} | {100 * ms / tot:.1f} % |")
    open(os.path.join(OUT, "r1_launches.md"), "w").write("\n".join(lines) + "\n")
    subprocess.run(["cp", os.path.join(SRC, "launches_r1.csv"), os.path.join(OUT, "r1_launches.csv")], check=True)


def full(rep, tag, title, n_rows, write_traffic):
    raw = subprocess.run(["ncu", "-i", os.path.join(SRC, rep), "--page", "raw", "--csv"], capture_output=True,
                         text=True).stdout
    open(os.path.join(OUT, f"r1_{tag}_raw.csv"), "w").write(raw)
    rows = list(csv.reader(raw.splitlines()))
    m = {h: (rows[2][i], rows[1][i]) for i, h in enumerate(rows[0])}
    out = [f"# ncu --set full, {title} (round 1)", "",
           f"`ncu --set full --clock-control none --import-source on -k regex:hfe_s -s 2 -c 1 python "
           f"tools/ncu_target.py` ({n_rows} synthetic rows; raw metric dump: `r1_{tag}_raw.csv`).", "",
           "| metric | value | unit |", "|---|---:|---|"]
    out += [f"| `{k}` | {m[k][0]} | {m[k][1]} |" for k in KEYS if k in m]
    tb = sum(float(m[k][0].replace(",", "")) * SCALE[m[k][1]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    out += ["", f"DRAM traffic per launch: {tb / 1e6:.1f} MB = **{tb / n_rows:.1f} B/row** against 76 B/row algorithmic "
                f"(68 read + 8 written).", ""]
    if write_traffic:
        path = os.path.join(OUT, "traffic.json")
        t = json.load(open(path)) if os.path.exists(path) else {}
        t[write_traffic] = tb / n_rows
        t[write_traffic + "_source"] = (f"profiles/r1_{tag}_raw.csv ({n_rows} rows, ncu --set full, dram__bytes_read.sum "
                                        f"+ dram__bytes_write.sum)")
        json.dump(t, open(path, "w"), indent=1)
    src = subprocess.run(["ncu", "-i", os.path.join(SRC, rep), "--page", "source", "--csv"], capture_output=True,
                         text=True).stdout
    srows = list(csv.reader(src.splitlines()))
    ci = {h: i for i, h in enumerate(srows[1])}
    data = srows[2:]
    samples = [int(r[ci["# Samples"]] or 0) for r in data]
    ex = [int(r[ci["Instructions Executed"]] or 0) for r in data]
    hot = [i for i, e in enumerate(ex) if e >= max(ex) * 0.9]
    tot = sum(samples)
    if tag == "fused" and len(hot) > 100:
        out += [f"Warp-state samples: RBF leaf loop ({len(hot)} SASS instructions per 4 support vectors) "
                f"{100 * sum(samples[i] for i in hot) / tot:.1f} %, before it {100 * sum(samples[:hot[0]]) / tot:.1f} %, "
                f"after it {100 * sum(samples[hot[-1] + 1:]) / tot:.1f} % (with the prefix-shared table layout the "
                f"prefix-quad steps, executed once per quad, count as 'before').", ""]
    open(os.path.join(OUT, f"r1_{tag}_kernel.md"), "w").write("\n".join(out) + "\n")


if __name__ == "__main__":
    launches()
    full("prof_r1_fused.ncu-rep", "fused", "fused kernel `hfe_score_kernel<SVC, stumps>`, published model (434 support "
         "vectors)", 4_000_000, "dram_bytes_per_row")
    full("prof_r1_nosvc.ncu-rep", "nosvc", "streaming pass `hfe_stream_kernel<stumps>` (no SVC member: the HBM-bound "
         "configuration)", 10_000_000, "dram_bytes_per_row_stream")
    if os.path.exists(os.path.join(SRC, "prof_stress.ncu-rep")):
        full("prof_stress.ncu-rep", "stress", "fused kernel `hfe_score_kernel<no SVC, general trees>` on the 2000-tree "
             "stress forest (BASELINE config 4)", 1_000_000, None)
    for name in ("retrained", "published", "nosvc", "stress", "reference"):
        p = os.path.join(SRC, f"r1_bench_{name}.json")
        if os.path.exists(p):
            open(os.path.join(OUT, f"r1_bench_{name}.json"), "w").write(open(p).read().strip().splitlines()[-1] + "\n")
    for name in ("scale2", "scale8", "scale8_1b", "scale8_nosvc", "scale8_stress", "perm8"):
        p = os.path.join(SRC, f"{name}.json")
        if os.path.exists(p):
            lines = [ln for ln in open(p).read().strip().splitlines() if ln.startswith("{")]
            keep = lines if name == "perm8" else lines[-1:]          # perm8: one line per shuffle mode
            open(os.path.join(OUT, f"r1_{name}.json"), "w").write("\n".join(keep) + "\n")
