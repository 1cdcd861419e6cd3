"""Summarise an .ncu-rep: one line per captured launch with the counters DESIGN.md quotes.
usage: python tools/ncu_summary.py report.ncu-rep > profiles/rNN_<what>_ncu_summary.txt"""
import csv
import io
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "us", 1e-3),
    ("dram__bytes_read.sum", "dram_rd_MB", 1e-6),
    ("dram__bytes_write.sum", "dram_wr_MB", 1e-6),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%", 1),
    ("lts__t_sector_hit_rate.pct", "l2_hit_%", 1),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_%", 1),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_%", 1),
    ("smsp__inst_executed.sum", "warp_instr", 1),
    ("launch__registers_per_thread", "regs", 1),
    ("launch__grid_size", "grid", 1),
    ("launch__block_size", "block", 1),
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv", "--print-units", "base"],
                         capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    print(f"# {path}: ncu --set full --clock-control none; per-launch values (cold caches, serialised)")
    print("kernel".ljust(34) + "".join(n.rjust(16) for _, n, _ in METRICS))
    for r in rows[2:]:
        name = r[ix["Kernel Name"]].split("(")[0].replace("klm::", "")
        cells = []
        for m, _, scale in METRICS:
            v = r[ix[m]] if m in ix else ""
            try:
                cells.append(f"{float(v.replace(',', '')) * scale:.2f}".rjust(16))
            except ValueError:
                cells.append("-".rjust(16))
        print(name[:33].ljust(34) + "".join(cells))


if __name__ == "__main__":
    main(sys.argv[1])
