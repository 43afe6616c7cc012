"""Turns gpurun_out/{launches_*.csv, prof_*.ncu-rep} into the tracked summaries under profiles/.
usage: python profiles/summarize.py <tag> <launches.csv> <prof.ncu-rep>"""
import collections
import csv
import subprocess
import sys

tag, launches, rep = sys.argv[1:4]
out = [f"# ncu summary {tag}", ""]
rows = list(csv.reader(open(launches, errors="ignore")))
hdr, data = None, []
for r in rows:
    if r and r[0] == "ID":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        data.append(dict(zip(hdr, r)))
agg = collections.defaultdict(lambda: [0, 0.0])
for d in data:
    v = float(d["Metric Value"].replace(",", ""))
    u = d["Metric Unit"]
    agg[d["Kernel Name"].split("(")[0][-48:]][0] += 1
    agg[d["Kernel Name"].split("(")[0][-48:]][1] += v * (1e-3 if u.startswith("ns") else 1.0 if u.startswith("us") else 1e3)
tot = sum(v for _, v in agg.values())
out += ["## launch list (gpu__time_duration.sum, --clock-control none; cold-cache, serialised: compare shares)", "",
        "| kernel | launches | total us | share |", "|---|---:|---:|---:|"]
for k, (n, v) in sorted(agg.items(), key=lambda x: -x[1][1])[:16]:
    out.append(f"| {k} | {n} | {v:.1f} | {v / tot:.3f} |")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = {n: i for i, n in enumerate(rows[0])}
cols = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp-inst")]
out += ["", "## --set full captures (per launch)", "", "| kernel | " + " | ".join(c[1] for c in cols) + " |",
        "|---|" + "---:|" * len(cols)]
for r in rows[2:]:
    name = r[h["Kernel Name"]].split("(")[0][-40:]
    vals = []
    for c, _ in cols:
        v = r[h[c]] if c in h else ""
        try:
            v = f"{float(v):.4g}"
        except ValueError:
            pass
        vals.append(v)
    out.append(f"| {name} | " + " | ".join(vals) + " |")
print("\n".join(out))
