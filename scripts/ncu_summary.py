"""Print the handful of ncu raw metrics used in profiles/README.md plus an instruction histogram by SASS region:
python scripts/ncu_summary.py report.ncu-rep [region_size]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
blk = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, r = rows[0], rows[-1]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"]
for i, k in enumerate(h):
    if k in keys or ("smsp__average_warps_issue_stalled" in k and "per_issue_active" in k and float(r[i] or 0) > 0.2):
        print(k, r[i])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = rows[1]
ia, isrc, isamp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
data = [(x[isrc].strip(), int(x[ia]), int(x[isamp])) for x in rows[2:] if len(x) > ia and x[ia].isdigit()]
tot = sum(d[1] for d in data)
tots = sum(d[2] for d in data)
print("total warp instructions", tot, "SASS lines", len(data))
for i in range(0, len(data), blk):
    b = data[i:i + blk]
    s, sm = sum(d[1] for d in b), sum(d[2] for d in b)
    if s > tot * 0.02 or sm > tots * 0.03:
        print(f"{i:5d} inst {100 * s / tot:5.1f}%  samples {100 * sm / max(tots, 1):5.1f}%  {b[0][0][:50]}")
