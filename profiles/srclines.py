#!/usr/bin/env python3
"""Per-source-line stall samples / instruction counts of one kernel from an .ncu-rep (needs -lineinfo).
usage: python profiles/srclines.py <report.ncu-rep> <kernel regex> [top N]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                      "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
cur, data, H = "", [], None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        H = r
        sa = H.index("Warp Stall Sampling (All Samples)"); ie = H.index("Instructions Executed")
    elif H and len(r) > ie and r[0] not in ("", "Function Name"):
        try:
            data.append((cur, r[0], r[1], int(r[sa] or 0), int(r[ie] or 0)))
        except ValueError:
            pass
tot = sum(d[3] for d in data) or 1; toti = sum(d[4] for d in data) or 1
print(f"kernel {kern}: {tot} stall samples, {toti} warp instructions")
for f, l, s, a, b in sorted(data, key=lambda x: -x[3])[:top]:
    print(f"{100*a/tot:5.1f}% samp {100*b/toti:5.1f}% inst | {f}:{l:>4s} {s.strip()[:105]}")
