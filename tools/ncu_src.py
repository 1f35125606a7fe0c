#!/usr/bin/env python
"""Summarise `ncu -i x.ncu-rep --page source --csv` (SASS view): stall-reason totals and the hottest instructions of
the k-th profiled kernel.  usage: ncu_src.py src.csv [k=0] [top=40]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
k = int(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
s = starts[k]
e = starts[k + 1] - 1 if k + 1 < len(starts) else len(rows)
hdr = rows[s]
body = [r for r in rows[s + 1:e] if len(r) >= len(hdr) - 2 and r[0].startswith("0x")]
ci = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
samp = lambda r: int(r[ci["# Samples"]] or 0)
tot = sum(samp(r) for r in body)
print(f"kernel {k}: {len(body)} SASS instructions, {tot} samples, "
      f"{sum(int(r[ci['Instructions Executed']] or 0) for r in body)} warp-instructions executed")
agg = {h: sum(int(r[ci[h]] or 0) for r in body) for h in stall_cols}
print("stall totals:", ", ".join(f"{h[6:]} {v} ({v/max(tot,1)*100:.0f}%)" for h, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
print(f"{'idx':>5s} {'samples':>7s} {'%':>5s} {'execs':>8s}  top stall           instruction")
order = sorted(range(len(body)), key=lambda i: -samp(body[i]))[:top]
for i in sorted(order):
    r = body[i]
    st = max(stall_cols, key=lambda h: int(r[ci[h]] or 0))
    print(f"{i:5d} {samp(r):7d} {samp(r)/max(tot,1)*100:5.1f} {r[ci['Instructions Executed']]:>8s}  {st[6:]:18s} {r[ci['Source']].strip()[:90]}")
