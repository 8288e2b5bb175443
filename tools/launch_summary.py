#!/usr/bin/env python3
"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv, collections, re, sys
rows = list(csv.reader(open(sys.argv[1])))
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0   # number of steps in the capture
hdr = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
H = rows[hdr]; ki = H.index('Kernel Name'); vi = H.index('Metric Value')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= vi: continue
    try: v = float(r[vi].replace(',', ''))
    except ValueError: continue
    n = re.sub(r'\(.*', '', r[ki])[:72]
    agg[n][0] += 1; agg[n][1] += v
tot = sum(v[1] for v in agg.values())
for n, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 24]:
    print(f"{v / 1e6 / div:9.3f} ms {100 * v / tot:5.1f}% x{c / div:6.1f} {n}")
print(f"total {tot / 1e6 / div:.3f} ms per step")
