#!/usr/bin/env python
"""Top source lines of one ncu capture by executed warp instructions / stall samples:
   python scripts/ncu_lines.py gpurun_out/prof.ncu-rep [N]"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = next(r for r in rows if r and r[0] == "Line No")
ix, sx = hdr.index("Instructions Executed"), hdr.index("# Samples")
lines = [r for r in rows if len(r) == len(hdr) and r[0].isdigit()]
tot = sum(int(r[ix] or 0) for r in lines)
ts = sum(int(r[sx] or 0) for r in lines)
print("warp instructions", tot, "samples", ts)
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
for r in sorted(lines, key=lambda r: -int(r[ix] or 0))[:top]:
    print(f"{r[0]:>5} {int(r[ix]):>10} {100*int(r[ix])/tot:5.1f}%  smp {100*int(r[sx] or 0)/max(ts,1):5.1f}%  {r[1].strip()[:120]}")
