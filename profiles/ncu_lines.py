#!/usr/bin/env python3
"""profiles/ncu_lines.py <csv from `ncu --page source --csv --print-source cuda,sass`> [top] — instructions executed and
stall samples aggregated per CUDA source line (needs -lineinfo)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, out, hdr = None, [], None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        iI, iS = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
        continue
    if cur and hdr and len(r) > 10 and r[0].isdigit() and r[2] == "-":
        try:
            out.append((cur, int(r[0]), int(r[iI]), int(r[iS]), r[1].strip()[:105]))
        except ValueError:
            pass
tot, tots = sum(o[2] for o in out), sum(o[3] for o in out)
print("total warp instructions", tot, "stall samples", tots)
for o in sorted(out, key=lambda o: -o[2])[:top]:
    print(f"{o[0]}:{o[1]:>4} inst {100 * o[2] / tot:5.1f}% stall {100 * o[3] / max(tots, 1):5.1f}% | {o[4]}")
