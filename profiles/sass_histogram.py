"""profiles/sass_histogram.py [lib.so] -> per-kernel counts of the SASS opcodes that show what the kernels are
built from: UBLKCP (TMA bulk copy), SYNCS (mbarrier), LDGSTS (cp.async), REDG / ATOMG (global atomics), ATOMS
(shared atomics), MATCH, MUFU, FFMA, and any tensor-core opcode (none expected on this path: UTC*MMA / HMMA)."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "eags-slam_b200/eags_raster/libeags_raster.so"
WANT = ("UBLKCP", "UTMALDG", "SYNCS", "LDGSTS", "REDG", "ATOMG", "ATOMS", "RED", "ATOM", "MATCH", "MUFU", "FFMA",
        "FMUL", "FADD", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "UTCHMMA", "UTCIMMA", "UTCQMMA", "HMMA", "IMMA")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn, cnt, tot = None, collections.defaultdict(collections.Counter), collections.Counter()
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        fn = re.sub(r"^_ZN\d+_GLOBAL__N__[0-9a-f]+_\d+_(\w+?)_cu_[0-9a-f]{8}\d+", r"\1::", m.group(1))
        fn = re.sub(r"^_Z\d+", "", fn)[:64]
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+)", ln)
    if m and fn:
        op = m.group(1)
        tot[fn] += 1
        if op in WANT:
            cnt[fn][op] += 1
print(f"# {lib}: SASS opcode histogram per kernel (cuobjdump -sass); tensor-core opcodes (UTC*MMA/HMMA/IMMA) expected absent")
for f in sorted(tot):
    print(f"{f:62s} total {tot[f]:6d} | " + " ".join(f"{k} {v}" for k, v in sorted(cnt[f].items())))
allops = collections.Counter()
for c in cnt.values():
    allops.update(c)
print("ALL:", " ".join(f"{k} {v}" for k, v in sorted(allops.items())))
