#!/usr/bin/env python3
"""One line per captured kernel of an .ncu-rep (ncu --set full): time, DRAM traffic, throughput shares, stalls.
usage: python profiles/ncu_table.py <report.ncu-rep> [more reports ...]"""
import csv
import io
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "inst"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg"),
        ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "st_math"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait")]
for rep in sys.argv[1:]:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    H, U = rows[0], rows[1]
    print(f"# {rep}: ncu --set full --clock-control none (cold-cache, serialised launches)")
    print("kernel".ljust(22) + "".join(n.rjust(10) for _, n in COLS))
    for r in rows[2:]:
        name = r[H.index("Kernel Name")].split("(")[0].split("::")[-1]
        line = name.ljust(22)
        for m, _ in COLS:
            if m not in H:
                line += "-".rjust(10)
                continue
            v, u = r[H.index(m)], U[H.index(m)]
            try:
                x = float(v.replace(",", ""))
                if u == "ns": x /= 1e3
                if u == "ms": x *= 1e3
                if u in ("byte", "Kbyte", "Mbyte", "Gbyte"):
                    x *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}[u]   # MB
                line += (f"{x:.0f}" if x >= 100 else f"{x:.2f}").rjust(10)
            except ValueError:
                line += v[:9].rjust(10)
        print(line)
    print("# rd / wr in MB per launch, us per launch, stalls = warps stalled per issue-active cycle")
