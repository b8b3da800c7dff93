#!/usr/bin/env python
"""Per CUDA source line: executed warp instructions and stall samples, from
`ncu -i rep --page source --csv --print-source cuda,sass`.  usage: ncu_line_summary.py file.csv [top]"""
import csv, sys
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None; out = []
for r in csv.reader(open(path)):
    if len(r) == 2 and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0].isdigit():
        try: out.append((cur_file, int(r[0]), r[1].strip(), float(r[6] or 0), float(r[7] or 0)))
        except ValueError: pass
ti = sum(o[4] for o in out); ts = sum(o[3] for o in out)
print(f"total instr {ti:.0f} samples {ts:.0f}")
for f, ln, src, s, n in sorted(out, key=lambda o: -o[4])[:top]:
    print(f"{f}:{ln:4d} instr {100*n/ti:5.1f}% samp {100*s/ts:5.1f}%  {src[:90]}")
