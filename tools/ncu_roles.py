"""Summarise an ncu capture of hfe_mma_kernel by warp role: executed instructions per tile, samples, barrier waits."""
import csv, subprocess, sys
rep, tiles = sys.argv[1], float(sys.argv[2])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, data = rows[1], rows[2:]
isrc, iex, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
st = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
idx = [hdr.index(h) for h in st]
def find(p): return [i for i, r in enumerate(data) if p in r[isrc]]
ldg, stg, ldtm = find("LDG"), find("STG"), find("LDTM")
print("n", len(data), "first LDG", ldg[:1], "STG", stg[:1], stg[-1:], "LDTM", ldtm, "UTCHMMA", find("UTCHMMA")[:1])
print("total exec/tile", sum(int(r[iex]) for r in data) / tiles, "samples", sum(int(r[ismp]) for r in data))
for i in [i for i, r in enumerate(data) if "SYNCS" in r[isrc] and "TRYWAIT" in r[isrc]]:
    sm = sum(int(r[ismp]) for r in data[i - 2:i + 22])
    print(f"  wait@{i}: spins/tile={int(data[i][iex]) / tiles:7.1f} samples~{sm:6d} {data[i][isrc].strip()[:64]}")
bounds = [int(a) for a in sys.argv[3:]] or [0, len(data)]
for lo, hi in zip(bounds[:-1], bounds[1:]):
    tot = {h: 0 for h in st}; ex = sm = 0
    for r in data[lo:hi]:
        ex += int(r[iex]); sm += int(r[ismp])
        for h, i in zip(st, idx):
            if r[i]: tot[h] += int(r[i])
    print(f"[{lo},{hi}) exec/tile={ex / tiles:8.1f} samples={sm}", {k.replace("stall_", ""): v for k, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v > 0.04 * max(sm, 1)})
