"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total, share."""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hdr]
ki, vi, ii = h.index("Kernel Name"), h.index("Metric Value"), h.index("ID")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[hdr + 1:]:
    try:
        if int(r[ii]) < skip:
            continue
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    name = r[ki].replace("lifter::<unnamed>::", "").replace("void ", "")[:70]
    tot[name] += v
    cnt[name] += 1
s = sum(tot.values())
print(f"total device time {s/1e3:.1f} us over {sum(cnt.values())} launches")
for k, v in sorted(tot.items(), key=lambda x: -x[1]):
    print(f"{k:72s} n={cnt[k]:4d} total={v/1e3:9.1f} us share={v/s*100:5.1f}% avg={v/cnt[k]/1e3:8.2f} us")
