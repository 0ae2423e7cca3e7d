"""Aggregates an ncu report's source page by CUDA source line: python scratch/ncu_src.py report.ncu-rep [launch_index] [top]"""
import csv, subprocess, sys
rep = sys.argv[1]; skip = sys.argv[2] if len(sys.argv) > 2 else "0"; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = None; hdr = None; agg = []
for r in rows:
    if r and r[0] == 'File Path': cur = r[1]; continue
    if r and r[0] == 'Function Name': print(r[1][:150]); continue
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr is None or len(r) < 8 or r[0] == '': continue
    try: s = int(r[6] or 0); ie = int(r[7] or 0)
    except ValueError: continue
    agg.append((s, ie, cur.split('/')[-1], r[0], r[1][:120]))
tot = sum(a[0] for a in agg) or 1; toti = sum(a[1] for a in agg) or 1
print('total samples', tot, 'instr', toti)
for a in sorted(agg, reverse=True)[:top]:
    print(f"{a[0]:6d} {100*a[0]/tot:5.1f}% inst={a[1]:8d} {100*a[1]/toti:4.1f}% {a[2]}:{a[3]}: {a[4]}")
