"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel time and share
of the last iteration (cold-cache, serialised -- compare shares, not absolutes)."""
import collections
import csv
import sys


def main(path, iters=3):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr, rows = rows[0], rows[1:]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    last = rows[(iters - 1) * len(rows) // iters:]
    per, tot = collections.OrderedDict(), 0.0
    for r in last:
        name = r[ki].split("(")[0].replace("void ", "").replace("<unnamed>::", "")[-70:]
        v = float(r[vi].replace(",", ""))
        v = {"ns": v / 1e3, "us": v, "ms": v * 1e3}.get(r[ui], v)
        per.setdefault(name, [0.0, 0])
        per[name][0] += v
        per[name][1] += 1
        tot += v
    print(f"{path}: {len(last)} launches in the last of {iters} iterations, {tot:.1f} us of kernel time")
    for k, (v, c) in per.items():
        print(f"  {v:9.1f} us  x{c}  {100 * v / tot:5.1f} %  {k}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 3)
