"""Per-kernel summary table of an ncu --set full report (the numbers DESIGN.md / bench.py quote).
Usage: python tools/ncu_summary.py REPORT.ncu-rep > profiles/<name>.txt"""
import csv
import io
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "dur"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("smsp__inst_executed.sum", "warp_inst")]


def main():
    raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    idx = [(hdr.index(c), n) for c, n in COLS if c in hdr]
    print("# %s  (ncu --set full --clock-control none; cold-cache, serialised launches)" % sys.argv[1])
    print("%-32s " % "kernel" + " ".join("%12s" % ("%s[%s]" % (n, units[i][:6]) if units[i] else n) for i, n in idx))
    for r in body:
        name = r[ki].split("(")[0][:32]
        out = []
        for i, n in idx:
            try:
                v = float(r[i])
                out.append("%12.4g" % v)
            except ValueError:
                out.append("%12s" % r[i][:12])
        print("%-32s " % name + " ".join(out))


if __name__ == "__main__":
    main()
