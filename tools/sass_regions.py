#!/usr/bin/env python
"""Aggregates an `ncu --page source --csv` export (SASS view) over windows of instructions: executed warp instructions, average
active lanes, stall samples and the dominant opcodes per window.  usage: python tools/sass_regions.py source.csv [window=100]"""
import csv, sys, collections
path = sys.argv[1]; win = int(sys.argv[2]) if len(sys.argv) > 2 else 100
rows = list(csv.reader(open(path)))
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0                       # which captured launch of the file
his = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hi = his[which]; end = his[which + 1] - 1 if which + 1 < len(his) else len(rows)
print(rows[hi - 1][1][:120] if hi > 0 else "")
hdr = rows[hi]; data = [r for r in rows[hi + 1:end] if len(r) == len(hdr)]
ci = {n: hdr.index(n) for n in ("Source", "# Samples", "Instructions Executed", "Thread Instructions Executed", "Warp Stall Sampling (All Samples)")}
tot_i = sum(float(r[ci["Instructions Executed"]]) for r in data); tot_t = sum(float(r[ci["Thread Instructions Executed"]]) for r in data)
tot_s = sum(float(r[ci["# Samples"]]) for r in data)
print(f"{len(data)} SASS instructions, {tot_i:.3g} warp instructions executed, {tot_t / max(tot_i, 1):.2f} lanes on average, {tot_s:.0f} samples")
for a in range(0, len(data), win):
    blk = data[a:a + win]
    wi = sum(float(r[ci["Instructions Executed"]]) for r in blk); ti = sum(float(r[ci["Thread Instructions Executed"]]) for r in blk)
    sm = sum(float(r[ci["# Samples"]]) for r in blk)
    ops = collections.Counter()
    for r in blk:
        ops[r[ci["Source"]].split()[0 if not r[ci["Source"]].strip().startswith("@") else 1].split(".")[0]] += float(r[ci["Instructions Executed"]])
    top = " ".join(f"{o}:{100 * n / max(wi, 1):.0f}" for o, n in ops.most_common(4))
    print(f"[{a:5d},{a + len(blk):5d}) exec {100 * wi / tot_i:5.1f}%  lanes {ti / max(wi, 1):5.1f}  samples {100 * sm / max(tot_s, 1):5.1f}%  {top}")
