#!/usr/bin/env python3
"""Dynamic SASS opcode histogram (per cell) + headline metrics of one kernel in an .ncu-rep.
usage: python profiles/opdist.py <report.ncu-rep> <kernel regex> [cells] [top]"""
import collections, csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
cells = float(sys.argv[3]) if len(sys.argv) > 3 else 8192.0 * 8192
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      f"regex:{kern}"], capture_output=True, text=True).stdout
H, cnt, tot = None, collections.Counter(), 0
for r in csv.reader(io.StringIO(raw)):
    if r and r[0] == "Address":
        H = r; si = H.index("Source"); ie = H.index("Instructions Executed"); continue
    if H and len(r) > ie:
        try: n = int(r[ie])
        except ValueError: continue
        op = r[si].split()
        if not op: continue
        o = op[1] if op[0].startswith("@") else op[0]
        cnt[".".join(o.split(".")[:3])] += n; tot += n
w = cells / 32
print(f"== {kern}: {tot / w:.1f} warp-instr per 32 cells")
for o, n in cnt.most_common(top): print(f"{o:28s} {n / w:7.2f} /cell  {100 * n / tot:5.1f}%")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
Hh, r = rows[0], rows[2]
for i, h in enumerate(Hh):
    if h in ("gpu__time_duration.sum", "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum",
             "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
             "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_active",
             "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
             "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
             "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "launch__grid_size", "launch__block_size") or \
       ("issue_stalled" in h and "per_issue_active" in h and float(r[i] or 0) > 0.1) or ("pipe_xu" in h):
        print(f"  {h:90s} {r[i]}")
