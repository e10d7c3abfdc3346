"""Turns an `ncu --metrics gpu__time_duration.sum --csv` log of bench.py into per-kernel shares of the LAST `--steps`
decode steps (the launch list committed under profiles/).  Per-launch times under ncu are cold-cache and serialised, so
the shares — not the absolutes — are what compares with the live bench.
Usage: python benchmarks/launch_shares.py gpurun_out/launches.csv [launches_per_step] [steps]"""
import csv
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    per_step = int(sys.argv[2]) if len(sys.argv) > 2 else 167
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            val = float(r["Metric Value"].replace(",", ""))
            unit = r["Metric Unit"]
            us = val / 1e3 if unit in ("ns", "nsecond") else (val if unit in ("us", "usecond") else val * 1e3)
            rows.append((r["Kernel Name"], r["Grid Size"], us))
    # the decode steps are the tail of the run; cut at the last `steps` sampling launches
    idx = [i for i, r in enumerate(rows) if "sample_kernel" in r[0]]
    if len(idx) >= steps + 1:
        lo, hi = idx[-steps - 1] + 1, idx[-1] + 1
        # the step's trailing bookkeeping (token feed, position increment) follows the sampling launch
        while hi < len(rows) and "ao::" not in rows[hi][0] and hi - idx[-1] <= 4:
            hi += 1
        sel = rows[lo:hi]
    else:
        sel = rows[-per_step * steps:]
    agg = OrderedDict()
    for name, grid, us in sel:
        short = name.replace("void ", "").split("(")[0][:70]
        k = (short, grid)
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    print(f"# {path}: {len(sel)} launches = {steps} decode steps, total {total:.1f} us under ncu (cold-cache, serialised)")
    for (name, grid), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{100 * us / total:6.2f}%  {us:9.1f} us  {n:4d} x  {name}  grid {grid}")


if __name__ == "__main__":
    main()
