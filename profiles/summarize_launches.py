"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals.
usage: python profiles/summarize_launches.py gpurun_out/launches.csv [skip_first_n] > profiles/rNN_launches.md"""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = []
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r["Metric Name"] == "gpu__time_duration.sum":
        rows.append((int(r["ID"]), r["Kernel Name"], float(r["Metric Value"])))
rows = [r for r in rows if r[0] >= skip]
agg = defaultdict(lambda: [0, 0.0])
for _, name, ns in rows:
    short = re.sub(r"\(.*", "", name)
    short = re.sub(r"^void ", "", short)
    agg[short][0] += 1
    agg[short][1] += ns
total = sum(v[1] for v in agg.values())
print(f"launches: {len(rows)}  total kernel time: {total / 1e6:.2f} ms (ncu: serialised, cold cache -- compare shares)\n")
print("| kernel | launches | total ms | share | ours |")
print("|---|---:|---:|---:|---|")
for name, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    ours = "yes" if name.startswith("iit::") else "torch"
    print(f"| `{name[:90]}` | {n} | {ns / 1e6:.3f} | {100 * ns / total:.1f}% | {ours} |")
