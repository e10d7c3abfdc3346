"""What the Blackwell-native claim rests on, as a committed listing: per kernel of lib/libao_b200.so, the counts of the
SASS mnemonics that identify the hardware paths (B200_PROFILING.md "What proves a Blackwell-native kernel"):

    UTC*MMA  tcgen05.mma          LDTM / STTM  tcgen05.ld / st (tensor memory)      UBLKCP  cp.async.bulk (1-D TMA)
    UTMALDG / UTMASTG  tensor-map TMA      SYNCS  mbarrier      HMMA / IMMA  mma.sync (legacy tensor path)
    LDSM  ldmatrix      MOVM  movmatrix      LDGSTS  cp.async

The last four columns are the kernel's static resources (`cuobjdump --dump-resource-usage`): registers per thread, stack
and local bytes per thread (local > 0 = spills or indexed private arrays), static shared bytes.

    python benchmarks/sass_census.py [lib] > profiles/r02_sass_census.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "torchao_fork_b200", "lib", "libao_b200.so")
PATTERNS = [("UTCMMA", r"\bUTC\w*MMA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UBLKCP", r"\bUBLKCP"),
            ("UTMALDG", r"\bUTMALDG"), ("UTMASTG", r"\bUTMASTG"), ("SYNCS", r"\bSYNCS"), ("HMMA", r"\bHMMA"),
            ("IMMA", r"\bIMMA"), ("LDSM", r"\bLDSM"), ("MOVM", r"\bMOVM"), ("LDGSTS", r"\bLDGSTS"), ("instr", r"^\s+/\*[0-9a-f]{4,}\*/\s+\S")]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
demangle = {}
names = sorted(set(re.findall(r"Function : (\S+)", out)))
if names:
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    demangle = dict(zip(names, dem))
counts, cur = collections.OrderedDict(), None
for line in out.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    for name, pat in PATTERNS:
        if re.search(pat, line):
            counts[cur][name] += 1
res = {}
ru = subprocess.run(["cuobjdump", "--dump-resource-usage", lib], capture_output=True, text=True).stdout.split("\n")
for i, line in enumerate(ru):
    m = re.search(r"Function (\S+):", line)
    if m and i + 1 < len(ru):
        res[m.group(1)] = {k: int(v) for k, v in re.findall(r"\b(REG|STACK|SHARED|LOCAL):(\d+)", ru[i + 1])}
cols = [n for n, _ in PATTERNS]
rcols = ["REG", "STACK", "LOCAL", "SHARED"]
print(f"# SASS census of {os.path.relpath(lib, ROOT)} (cuobjdump -sass, sm_100a); one row per kernel")
print("  ".join(f"{c:>7s}" for c in cols + rcols) + "  kernel")
tot = collections.Counter()
for fn, c in counts.items():
    tot.update(c)
    short = re.sub(r"\(.*", "", demangle.get(fn, fn))
    short = re.sub(r"^void ", "", short)
    r = res.get(fn, {})
    print("  ".join([f"{c[k]:7d}" for k in cols] + [f"{r.get(k, -1):7d}" for k in rcols]) + "  " + short)
print("  ".join(f"{tot[k]:7d}" for k in cols) + "  TOTAL")
