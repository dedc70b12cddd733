"""Summarise an `ncu --set full` report (read here, no GPU needed) as a markdown table for profiles/.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_xx_ncu_full.md
"""
import csv
import io
import subprocess
import sys

COLS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %"),
    ("launch__registers_per_thread", "regs"),
    ("smsp__inst_executed.sum", "warp insts"),
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    sel = [(hdr.index(m), n) for m, n in COLS if m in hdr]
    print("| kernel | " + " | ".join(n for _, n in sel) + " |")
    print("|---|" + "---|" * len(sel))
    for r in rows[2:]:
        name = r[ki].split("(")[0].replace("void ", "")
        cells = []
        for i, n in sel:
            v = r[i]
            try:
                f = float(v)
                v = ("%.0f" % f) if f >= 1000 or f == int(f) else ("%.3g" % f)
            except ValueError:
                pass
            u = units[i]
            cells.append(v + (" " + u if u in ("ms", "us", "Mbyte", "Gbyte", "Kbyte") else ""))
        print("| %s | %s |" % (name, " | ".join(cells)))


if __name__ == "__main__":
    main()
