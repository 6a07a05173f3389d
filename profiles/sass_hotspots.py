"""Group the SASS of one kernel (ncu --page source --csv) into runs of equal execution count and
print each run's share of executed warp instructions + opcode mix.
usage: ncu -i rep --page source --csv --kernel-name regex:NAME > x.csv; python sass_hotspots.py x.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
hdr = rows[h]
ie, isrc, ismp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
data = []
for r in rows[h + 1:]:
    if len(r) <= ie:
        continue
    try:
        data.append((r[isrc].strip(), int(r[ie]), int(r[ismp])))
    except ValueError:
        pass
tot = sum(d[1] for d in data)
print("kernel:", rows[0][1][:80] if len(rows[0]) > 1 else "?", "| warp instructions", tot, "| samples", sum(d[2] for d in data))


def op(s):
    t = s.split()
    t = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
    return t.split(".")[0]


groups = []
for s, e, m in data:
    if groups and abs(groups[-1]["e"] - e) <= 0.03 * max(e, 1):
        g = groups[-1]
        g["n"] += 1; g["sum"] += e; g["smp"] += m; g["ops"].append(op(s))
    else:
        groups.append(dict(e=e, n=1, sum=e, smp=m, ops=[op(s)]))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
for g in groups:
    if g["sum"] > thr * tot:
        c = collections.Counter(g["ops"])
        print(f"exec/inst {g['e']:>10d}  n_inst {g['n']:4d}  share {100 * g['sum'] / tot:5.1f}%  samples {g['smp']:6d}  ",
              dict(c.most_common(9)))
