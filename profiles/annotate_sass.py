"""Join an ncu source-page CSV (SASS view, per-instruction executed counts) with nvdisasm line info.
Usage: python profiles/annotate_sass.py <report.ncu-rep> <object.o> <kernel-name-substring> [warps]
Prints per-source-line executed warp instructions (per warp if `warps` is given) and writes /tmp/annot.txt."""
import collections
import csv
import re
import subprocess
import sys
import tempfile
import os

rep, obj, pat = sys.argv[1:4]
warps = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
txt = subprocess.check_output(["nvdisasm", "--print-line-info", os.path.join(tmp, cubin)], text=True).split("\n")
src = subprocess.check_output(["ncu", "-i", rep, "--page", "source", "--csv"], text=True)
rows = list(csv.reader(src.split("\n")))
kname = rows[0][1]
hdr = rows[1]
ie, sm = hdr.index("Instructions Executed"), hdr.index("# Samples")
data = [r for r in rows[2:] if len(r) > sm]
start = None
for i, l in enumerate(txt):
    if l.startswith(".text.") and pat in l:
        start = i
        break
seq, cur = [], None
for l in txt[start + 1:]:
    if (l.startswith(".text.") or l.startswith(".section")) and seq:
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        seq.append((int(m.group(1), 16), m.group(2), cur))
assert len(seq) == len(data), (len(seq), len(data), kname)
agg, smp = collections.Counter(), collections.Counter()
with open("/tmp/annot.txt", "w") as f:
    for (a, ins, c), r in zip(seq, data):
        n = int(r[ie])
        agg[c] += n
        smp[c] += int(r[sm])
        f.write("%04x %7.2f %4s  %-72s %s:%d\n" % (a, n / warps, r[sm], ins[:72], c[0] if c else "-", c[1] if c else 0))
tot, ts = sum(agg.values()), sum(smp.values())
print(kname)
print("total %.1f" % (tot / warps))
byfile = collections.Counter()
for k, v in agg.items():
    byfile[k[0] if k else None] += v
print({k: round(v / warps, 1) for k, v in byfile.items()})
for k, v in sorted(agg.items(), key=lambda kv: kv[0] or ("", 0)):
    if v / tot >= 0.004:
        print(k, "%.1f" % (v / warps), "%.1f%%" % (100 * smp[k] / ts))
