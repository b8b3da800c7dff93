#!/usr/bin/env python
"""Summarise `ncu --page source --csv`: executed instructions per opcode and the hottest
SASS lines by stall samples.  usage: ncu_source_summary.py file.csv [top]"""
import csv, sys, collections
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
recs = [r for r in rows[hi + 1:] if len(r) >= len(hdr) - 2]
def num(r, k):
    try: return float(r[ix[k]].replace(",", ""))
    except Exception: return 0.0
byop = collections.Counter(); samp = collections.Counter()
tot_i = tot_s = 0
for r in recs:
    src = r[ix["Source"]].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
    op = op.rstrip(";")
    n = num(r, "Instructions Executed"); s = num(r, "# Samples")
    byop[op] += n; samp[op] += s; tot_i += n; tot_s += s
print(f"total warp instructions {tot_i:.0f}, samples {tot_s:.0f}")
print("--- by opcode (instructions, share, samples share)")
for op, n in byop.most_common(30):
    print(f"{op:32s} {n:14.0f} {100*n/tot_i:6.2f}% {100*samp[op]/max(tot_s,1):6.2f}%")
print("--- hottest lines by samples")
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r in sorted(recs, key=lambda r: -num(r, "# Samples"))[:top]:
    st = sorted(((num(r, c), c) for c in stall_cols), reverse=True)[:2]
    print(f"{r[ix['Address']][-5:]} {num(r,'# Samples'):7.0f} {num(r,'Instructions Executed'):11.0f}  {r[ix['Source']].strip()[:70]:70s} " + " ".join(f"{c[6:]}={v:.0f}" for v, c in st if v))
