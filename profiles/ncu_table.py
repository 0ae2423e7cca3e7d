"""Per-kernel table from `ncu -i rep.ncu-rep --page raw --csv` of one profiled step: launches grouped by kernel name with
summed duration, DRAM bytes, achieved DRAM GB/s, tensor-pipe activity, registers, occupancy; also refreshes
profiles/ncu_traffic.json (dram bytes per launch by kernel name, read by bench.py's roofline.traffic).
usage: python profiles/ncu_table.py raw.csv out.md "title" [traffic-key-suffix]"""
import csv
import json
import os
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {k: i for i, k in enumerate(hdr)}


def num(r, key):
    if key not in idx or r[idx[key]] in ("", "n/a"):
        return 0.0
    v = float(r[idx[key]].replace(",", ""))
    u = units[idx[key]]
    scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "byte": 1, "usecond": 1e-6, "msecond": 1e-3, "second": 1, "nsecond": 1e-9,
             "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1}.get(u)
    return v * scale if scale is not None else v


def short(name):
    name = name.replace("void ", "").replace("iit::", "")
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("__nv_bfloat16", "bf16")
    return name[:90]


groups = {}
for r in rows[2:]:
    k = short(r[idx["Kernel Name"]])
    g = groups.setdefault(k, {"n": 0, "t": 0.0, "rd": 0.0, "wr": 0.0, "tensor": 0.0, "regs": 0, "occ": 0.0, "l2": 0.0, "dram_pct": 0.0})
    t = num(r, "gpu__time_duration.sum")
    g["n"] += 1
    g["t"] += t
    g["rd"] += num(r, "dram__bytes_read.sum")
    g["wr"] += num(r, "dram__bytes_write.sum")
    g["tensor"] += num(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active") * t
    g["dram_pct"] += num(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed") * t
    g["occ"] += num(r, "sm__warps_active.avg.pct_of_peak_sustained_active") * t
    g["regs"] = max(g["regs"], int(num(r, "launch__registers_per_thread")))
    g["l2"] += num(r, "l1tex__m_xbar2l1tex_read_bytes.sum")
total = sum(g["t"] for g in groups.values())
with open(sys.argv[2], "w") as f:
    f.write(f"{sys.argv[3]}\n\n`ncu --set full --clock-control none` (per-launch times are serialised and cold-cache: compare SHARES, not "
            f"absolutes). {sum(g['n'] for g in groups.values())} launches, {total * 1e3:.2f} ms of kernels.\n\n")
    f.write("| kernel | launches | total us | share | avg us | DRAM rd+wr MB/launch | DRAM GB/s | DRAM % of peak | tensor pipe active % | warps active % | regs |\n")
    f.write("|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|\n")
    for k, g in sorted(groups.items(), key=lambda kv: -kv[1]["t"]):
        by = g["rd"] + g["wr"]
        f.write(f"| `{k}` | {g['n']} | {g['t'] * 1e6:.0f} | {100 * g['t'] / total:.1f}% | {g['t'] / g['n'] * 1e6:.1f} | {by / g['n'] / 1e6:.1f} | "
                f"{by / g['t'] / 1e9 if g['t'] else 0:.0f} | {g['dram_pct'] / g['t'] if g['t'] else 0:.0f} | {g['tensor'] / g['t'] if g['t'] else 0:.1f} | "
                f"{g['occ'] / g['t'] if g['t'] else 0:.0f} | {g['regs']} |\n")
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_traffic.json")
table = json.load(open(path)) if os.path.exists(path) else {}
suffix = sys.argv[4] if len(sys.argv) > 4 else ""
for k, g in groups.items():
    base = k.split("<")[0] + suffix
    ent = table.setdefault(base, {"dram_bytes_per_launch": 0.0, "launches": 0, "source": sys.argv[3]})
    ent.update({"dram_bytes_per_launch": (g["rd"] + g["wr"]) / g["n"], "launches": g["n"], "source": sys.argv[3]})
json.dump(table, open(path, "w"), indent=1)
print(f"{len(groups)} kernels, {total * 1e3:.2f} ms")
