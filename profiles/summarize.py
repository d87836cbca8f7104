#!/usr/bin/env python3
"""Turn gpurun_out/{launches_*.csv, prof_*.ncu-rep} into the small text summaries kept under profiles/.
usage: python profiles/summarize.py <tag> [steps_in_capture]   (reads gpurun_out/launches_<tag>.csv, prof_<tag>.ncu-rep)"""
import collections
import csv
import io
import subprocess
import sys

tag = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
out = io.StringIO()

rows = list(csv.reader(open(f"gpurun_out/launches_{tag}.csv")))
h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H, data = rows[h], rows[h + 1:]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.OrderedDict()
for r in data:
    if len(r) <= vi:
        continue
    n = r[ki].split("(")[0][:70]
    f = {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[ui], 1e-6)
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += float(r[vi].replace(",", "")) * f
tot = sum(a[1] for a in agg.values())
out.write(f"# ncu launch list ({tag}): gpu__time_duration.sum, --clock-control none; {len(data)} launches, "
          f"{steps} composite steps + 1 DEM generation in the capture\n")
out.write("# cold-cache, serialised launches: compare SHARES, not absolutes\n")
out.write(f"{'kernel':70s} {'launches':>8s} {'total ms':>10s} {'share':>7s}\n")
for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    out.write(f"{n:70s} {a[0]:8d} {a[1]:10.3f} {100 * a[1] / tot:6.1f}%\n")
out.write(f"{'TOTAL':70s} {len(data):8d} {tot:10.3f}\n\n")

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__grid_size",
        "launch__block_size", "smsp__inst_executed.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_issue_stalled_barrier.pct", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]
try:
    raw = subprocess.run(["ncu", "-i", f"gpurun_out/prof_{tag}.ncu-rep", "--page", "raw", "--csv"],
                         capture_output=True, text=True, check=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    H = rr[0]
    out.write(f"# ncu --set full --clock-control none ({tag}): one launch per kernel\n")
    for r in rr[2:]:
        out.write(f"\n## {r[H.index('Kernel Name')][:110]}\n")
        for w in WANT:
            if w in H:
                out.write(f"  {w:82s} {r[H.index(w)]:>16s} {rr[1][H.index(w)]}\n")
        rd = float(r[H.index("dram__bytes_read.sum")].replace(",", ""))
        wr = float(r[H.index("dram__bytes_write.sum")].replace(",", ""))
        ur, uw = rr[1][H.index("dram__bytes_read.sum")], rr[1][H.index("dram__bytes_write.sum")]
        sc = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        out.write(f"  {'traffic = dram read + write':82s} {(rd * sc[ur] + wr * sc[uw]) / 1e6:16.1f} MB\n")
except Exception as e:  # noqa: BLE001
    out.write(f"# no full capture for {tag}: {e}\n")
open(f"profiles/ncu_{tag}.txt", "w").write(out.getvalue())
print(out.getvalue())
