"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (usage: python summarize_launches.py file.csv)."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
H = rows[hdr]
data = rows[hdr + 2:]
ki, vi = H.index('Kernel Name'), H.index('Metric Value')
agg = collections.OrderedDict()
for r in data:
    if len(r) <= vi:
        continue
    n = r[ki].split('(')[0][:60]
    agg.setdefault(n, []).append(float(r[vi].replace(',', '')))
tot = sum(sum(v) for v in agg.values())
for n, v in agg.items():
    print("%-62s n=%3d  avg %10.1f us  tot %10.1f us  %5.1f%%" % (n, len(v), sum(v) / len(v) / 1e3, sum(v) / 1e3, 100 * sum(v) / tot))
print("total ms", tot / 1e6)
