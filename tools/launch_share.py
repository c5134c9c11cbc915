#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.  usage: python tools/launch_share.py launches.csv"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]; k = hdr.index("Kernel Name"); v = hdr.index("Metric Value"); u = hdr.index("Metric Unit")
tot = collections.Counter(); cnt = collections.Counter()
for r in rows[hi + 1:]:
    if len(r) <= v: continue
    name = r[k].split("(")[0].replace("hrt::<unnamed>::", "").replace("void ", "")[-48:]
    t = float(r[v].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[u], 1.0)
    tot[name] += t; cnt[name] += 1
T = sum(tot.values())
print(f"{sum(cnt.values())} launches, {T / 1e3:.2f} ms of kernel time (serialised, cold caches)")
for n, t in tot.most_common(): print(f"{n:50s} {cnt[n]:6d} launches {t / 1e3:10.2f} ms {100 * t / T:5.1f} %  avg {t / cnt[n]:8.1f} us")
