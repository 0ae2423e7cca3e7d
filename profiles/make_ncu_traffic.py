"""profiles/ncu_traffic.json from an `ncu --set full ... --page raw --csv` dump of `bench.py --profile-forward`:
average DRAM traffic per launch of the kernels behind iit_conv2d_fprop_tc (one LL forward = its launches of a step).
usage: ncu -i rep.ncu-rep --page raw --csv > raw.csv; python profiles/make_ncu_traffic.py raw.csv [table.md]"""
import csv
import json
import os
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
col = {k: hdr.index(k) for k in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                 "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "launch__grid_size",
                                 "l1tex__m_xbar2l1tex_read_bytes.sum")}
units = rows[1]


def to_bytes(v, unit):
    return float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


launches = []
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    if "conv3x3_tc" not in name and "conv_win_kernel" not in name:
        continue
    rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
    wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
    xb = to_bytes(r[col["l1tex__m_xbar2l1tex_read_bytes.sum"]], units[col["l1tex__m_xbar2l1tex_read_bytes.sum"]])
    launches.append({"kernel": name.split("(")[0].replace("void iit::", ""), "grid": r[col["launch__grid_size"]],
                     "us": float(r[col["gpu__time_duration.sum"]]), "dram_read": rd, "dram_write": wr, "l2_to_sm": xb,
                     "tensor_pct": float(r[col["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]])})
total = sum(l["dram_read"] + l["dram_write"] for l in launches)
out = {"iit_conv2d_fprop_tc": {"dram_bytes_per_launch": total / max(len(launches), 1), "launches": len(launches),
                               "dram_bytes_per_forward": total,
                               "source": "ncu --set full --clock-control none, bench.py --profile-forward (one LL forward, batch 4096)"}}
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_traffic.json"), "w"), indent=1)
if len(sys.argv) > 2:
    with open(sys.argv[2], "w") as f:
        f.write("ncu --set full, one LL forward at batch 4096 (bf16): the launches behind iit_conv2d_fprop_tc / iit_stem_conv_s2d\n\n")
        f.write("| kernel | grid | us | DRAM read MB | DRAM write MB | L2->SM MB | tensor pipe active % |\n|---|---:|---:|---:|---:|---:|---:|\n")
        for l in launches:
            f.write(f"| `{l['kernel']}` | {l['grid']} | {l['us']:.1f} | {l['dram_read'] / 1e6:.1f} | {l['dram_write'] / 1e6:.1f} | "
                    f"{l['l2_to_sm'] / 1e6:.1f} | {l['tensor_pct']:.1f} |\n")
print(json.dumps(out))
