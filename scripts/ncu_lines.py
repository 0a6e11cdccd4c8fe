"""Instruction / stall-sample histogram by CUDA source line of one ncu report: python scripts/ncu_lines.py report.ncu-rep [min_pct]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if r and r[0] == "Line No")
ia, isamp = hdr.index("Instructions Executed"), hdr.index("# Samples")
data = [(int(r[0]), r[1], int(r[ia]), int(r[isamp]) if r[isamp].isdigit() else 0) for r in rows if r and r[0].isdigit() and len(r) > ia and r[ia].isdigit()]
tot, ts = sum(d[2] for d in data), sum(d[3] for d in data)
print("warp instructions", tot, "samples", ts)
for d in data:
    if d[2] > tot * minpct / 100 or d[3] > ts * minpct / 100:
        print(f"{d[0]:4d} inst {100 * d[2] / tot:5.1f}%  samples {100 * d[3] / max(ts, 1):5.1f}%  {d[1].strip()[:100]}")
