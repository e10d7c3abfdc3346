"""Per-source-line stall samples of one kernel from an `ncu --set full --import-source on` report, without the GUI:
joins `ncu --page source --csv` (per SASS instruction) with the line table of `nvdisasm -g` on a cubin of the same build.

    python benchmarks/ncu_source_lines.py report.ncu-rep torchao_fork_b200/csrc/decode_mega.cu 'decode_mega_kernelILi128E' [top]
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys

rep, src, pat = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
cubin = "/tmp/_ncu_lines.cubin"
subprocess.run(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
                "--expt-relaxed-constexpr", "-DAO_B200_BUILD", "-cubin", "-o", cubin, src], check=True)
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
# the kernel's section: from its .text header to the next .section
start = next(i for i, l in enumerate(dis) if l.startswith("\t.section\t.text.") and pat in l)
end = next((i for i in range(start + 1, len(dis)) if dis[i].startswith("\t.section")), len(dis))
off2line, line = {}, None
base_name = os.path.basename(src)
for l in dis[start:end]:
    m = re.search(r'//## File "([^"]*)", line (\d+)', l)
    if m:
        line = int(m.group(2)) if os.path.basename(m.group(1)) == base_name else -1
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", l)
    if m:
        off2line[int(m.group(1), 16)] = line
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[hi], rows[hi + 1:]
ia, isamp, iex = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]


def num(x):
    try:
        return int(x, 16) if x.startswith("0x") else int(x)
    except Exception:
        return None


base = num(data[0][ia])
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
tot = totex = 0
for r in data:
    a = num(r[ia]) if r else None
    if a is None:
        continue
    ln = off2line.get(a - base, -2)
    s, ex = int(r[isamp] or 0), int(r[iex] or 0)
    agg[ln][0] += s
    agg[ln][1] += ex
    tot += s
    totex += ex
    for i in stall_cols:
        if r[i] and r[i] != "0":
            try:
                agg[ln][2][hdr[i]] += int(r[i])
            except ValueError:
                pass
text = open(src).read().split("\n")
print(f"total samples {tot}, instructions {totex}")
for ln, (s, ex, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    code = text[ln - 1].strip()[:80] if ln > 0 else "(other file / inlined header)"
    print(f"{ln:5d} {100 * s / max(tot, 1):5.1f}% samp {100 * ex / max(totex, 1):5.1f}% inst {dict(st.most_common(2))} | {code}")
