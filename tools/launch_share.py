#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections, csv, sys
rows = [l for l in open(sys.argv[1]) if not l.startswith("==")]
r = list(csv.reader(rows)); h = r[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
agg = collections.OrderedDict()
for row in r[1 + skip:]:
    k = row[ki][:80]; a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += float(row[vi].replace(",", ""))
tot = sum(a[1] for a in agg.values())
print(f"total {tot/1e6:.3f} ms over {sum(a[0] for a in agg.values())} launches")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1])[:25]:
    print(f"{a[1]/1e6:9.3f} ms {a[0]:5d} {100*a[1]/tot:5.1f}%  {k}")
