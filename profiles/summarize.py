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

# The roofline object of bench.py reads the dominant kernel's DRAM traffic from here (never a literal in bench.py):
# dram__bytes_read.sum + dram__bytes_write.sum of one decode_tc_kernel launch in THIS capture.
import json
import os
units = rows[1]
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for r in rows[2:]:
    if "decode_tc_kernel" in r[h["Kernel Name"]]:
        tot_b = sum(float(r[h[c]].replace(",", "")) * scale.get(units[h[c]], 1.0)
                    for c in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "roofline_traffic.json")
        json.dump({"kernel": "decode_tc_kernel", "dram_bytes_per_launch": tot_b,
                   "source": f"profiles/{tag}_summary.md: ncu --set full --clock-control none, one launch (cold caches, serialised)",
                   "captured_at_head": head}, open(path, "w"), indent=1)
        break
