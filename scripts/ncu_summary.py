"""Summarise an `ncu --set full` report (read on the CPU box): one row per captured launch.
usage: python scripts/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.txt"""
import csv
import io
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
cols = [("Kernel Name", "kernel", 44), ("launch__grid_size", "grid", 6), ("launch__block_size", "blk", 4),
        ("launch__registers_per_thread", "regs", 4), ("gpu__time_duration.sum", "us", 8),
        ("dram__bytes_read.sum", "rd_MB", 8), ("dram__bytes_write.sum", "wr_MB", 8),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 6),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%", 7),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 6),
        ("smsp__inst_executed.sum", "warp_inst", 10),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 5)]
idx = [(hdr.index(c), n, w) for c, n, w in cols if c in hdr]
print("  ".join(n.ljust(w) for _, n, w in idx))
for r in data:
    out = []
    for i, n, w in idx:
        v = r[i]
        if n == "kernel":
            v = v.replace("void ", "").replace("ditsep::", "").split("(CUtensor")[0].split("(const")[0].split("(Stats")[0]
        else:
            try:
                f = float(v)
                if n in ("rd_MB", "wr_MB") and units[i].lower().startswith("kbyte"):
                    f /= 1e3
                if n in ("rd_MB", "wr_MB") and units[i].lower() == "byte":
                    f /= 1e6
                v = f"{f:.0f}" if f >= 1000 or f == int(f) else f"{f:.2f}"
            except ValueError:
                pass
        out.append(v[:w].ljust(w))
    print("  ".join(out))
