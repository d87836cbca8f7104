#!/usr/bin/env python3
"""Per-source-line instruction / stall shares of kernels in an .ncu-rep (needs -lineinfo).
usage: python profiles/instdist.py <report.ncu-rep> <kernel regex> [top N] [cells]"""
import collections, csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
cells = float(sys.argv[4]) if len(sys.argv) > 4 else 8192.0 * 8192
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      f"regex:{kern}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
cur, H, data = "", None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        H = r; sa = H.index("Warp Stall Sampling (All Samples)"); ie = H.index("Instructions Executed")
    elif H and len(r) > ie and r[0] not in ("", "Function Name"):
        try: data.append((cur, int(r[0]), r[1], int(r[sa] or 0), int(r[ie] or 0)))
        except ValueError: pass
toti = sum(d[4] for d in data) or 1; tots = sum(d[3] for d in data) or 1
print(f"== {kern}: {toti/1e6:.0f}M warp instr, {toti*32/cells:.0f} thread-instr per cell")
agg = collections.OrderedDict()
for f, l, s, a, b in data:
    agg.setdefault((f, l), [s, 0, 0]); agg[(f, l)][1] += a; agg[(f, l)][2] += b
for (f, l), (s, a, b) in sorted(agg.items(), key=lambda x: -x[1][2])[:top]:
    print(f"{100*b/toti:5.1f}% inst {100*a/tots:5.1f}% samp | {f}:{l:4d} {s.strip()[:105]}")
