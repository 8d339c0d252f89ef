"""Summarise an .ncu-rep (read here, no GPU needed) into a markdown table for profiles/."""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]


def main(rep, title):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    cols = [c for c in WANT if c in hdr]
    print(f"## {title}\n\nsource: `{rep}` (`ncu --set full --clock-control none --import-source on`)\n")
    print("| kernel | grid | " + " | ".join(cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "")
        vals = [f"{r[hdr.index(c)]} {units[hdr.index(c)]}".strip() for c in cols]
        print(f"| {name} | {r[hdr.index('Grid Size')]} | " + " | ".join(vals) + " |")
    print()


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
