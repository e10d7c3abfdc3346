"""Turns an ncu report (.ncu-rep) into the compact per-launch table committed under profiles/.
Usage: python benchmarks/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.txt"""
import csv
import io
import subprocess
import sys

COLS = [
    ("Kernel Name", "kernel", 34), ("Grid Size", "grid", 12), ("Block Size", "block", 12),
    ("gpu__time_duration.sum", "time", 9), ("dram__bytes_read.sum", "dram_rd", 10), ("dram__bytes_write.sum", "dram_wr", 9),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 7),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%", 6),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%", 6),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%", 8),
    ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue%", 7),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 6),
    ("launch__registers_per_thread", "regs", 5),
    ("launch__shared_mem_per_block_dynamic", "smem_dyn", 9),
    ("smsp__inst_executed.sum", "warp_inst", 11),
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    idx = [(hdr.index(c) if c in hdr else None, name, w) for c, name, w in COLS]
    print(f"# {rep}: ncu --set full --clock-control none (cold-cache, serialised launches; compare shares, not absolutes)")
    print(" ".join(f"{name + ('[' + units[i] + ']' if i is not None and units[i] else ''):>{w}}" for i, name, w in idx))
    for r in body:
        out = []
        for i, name, w in idx:
            v = "-" if i is None else r[i]
            if name == "kernel":
                v = v.replace("void ", "").split("(")[0][:w]
            else:
                try:
                    v = f"{float(v):.2f}" if "." in v else v
                except ValueError:
                    pass
            out.append(f"{v:>{w}}")
        print(" ".join(out))


if __name__ == "__main__":
    main()
