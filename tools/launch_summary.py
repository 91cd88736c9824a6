"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel (development aid)."""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
tot = defaultdict(lambda: [0, 0.0])
for r in rows:
    name, unit, val = r[4], r[13], float(r[14].replace(",", ""))
    ms = val / 1e6 if unit in ("ns", "nsecond") else val / 1e3 if unit in ("us", "usecond") else val
    tot[name][0] += 1
    tot[name][1] += ms
ours = sum(v[1] for k, v in tot.items() if "b200m::" in k)
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1])[:16]:
    mine = "b200m::" in k
    print(f"{('' if mine else '(torch) ') + k[:100]:110s} launches={v[0]:5d} total_ms={v[1]:10.3f} share_of_b200m={100 * v[1] / ours if mine else 0:6.2f}%")
