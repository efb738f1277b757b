"""Summarise an `ncu --metrics ... --csv` launch list per kernel (usage: python summarize_launches.py file.csv).
Handles one or several metrics per launch; the duration metric gives avg / total / share of the listed launches,
DRAM bytes (if collected) give per-launch traffic and GB/s."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
H = rows[hdr]
ki, mi, ui, vi, gi = (H.index(c) for c in ('Kernel Name', 'Metric Name', 'Metric Unit', 'Metric Value', 'Grid Size'))
launch = collections.OrderedDict()          # ID -> {metric: value}
for r in rows[hdr + 1:]:
    if len(r) <= vi or not r[0].isdigit():
        continue
    d = launch.setdefault(r[0], {"name": r[ki].split('(')[0].replace('void ', '')[:52], "grid": r[gi]})
    try:
        v = float(r[vi].replace(',', ''))
    except ValueError:
        continue
    unit = r[ui]
    if r[mi] == 'gpu__time_duration.sum':
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(unit, 1e-3)
    if r[mi].startswith('dram__bytes'):
        v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
    d[r[mi]] = v
agg = collections.OrderedDict()
for d in launch.values():
    agg.setdefault(d["name"], []).append(d)
tot = sum(d.get('gpu__time_duration.sum', 0.0) for d in launch.values())
print("%-52s %4s %10s %10s %6s %9s %8s %6s %5s" % ("kernel", "n", "avg us", "tot us", "share", "dram MB", "GB/s", "occ%", "regs"))
for n, ds in agg.items():
    t = [d.get('gpu__time_duration.sum', 0.0) for d in ds]
    by = [d.get('dram__bytes_read.sum', 0.0) + d.get('dram__bytes_write.sum', 0.0) for d in ds]
    occ = [d.get('sm__warps_active.avg.pct_of_peak_sustained_active', float('nan')) for d in ds]
    regs = ds[0].get('launch__registers_per_thread', float('nan'))
    at, ab = sum(t) / len(t), sum(by) / len(by)
    print("%-52s %4d %10.1f %10.1f %5.1f%% %9.1f %8.1f %6.1f %5.0f" % (n, len(ds), at, sum(t), 100 * sum(t) / tot if tot else 0,
                                                                  ab / 1e6, ab / at / 1e3 if at else 0, sum(occ) / len(occ), regs))
print("total listed: %.3f ms in %d launches (ncu serialises launches and runs them cold: shares, not absolutes, carry over)"
      % (tot / 1e3, len(launch)))
