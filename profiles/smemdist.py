#!/usr/bin/env python3
"""Shared-memory wavefronts per SASS instruction of one kernel in an .ncu-rep (source page), with the source line.
usage: python profiles/smemdist.py <report.ncu-rep> <kernel regex> [top]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      f"regex:{kern}"], capture_output=True, text=True).stdout
H = None
rows = []
ntab = 0
for r in csv.reader(io.StringIO(raw)):
    if r and r[0] == "Address":
        ntab += 1
        if ntab > 1: break
        H = r
        si, wi, ii, xi, ei = (H.index(k) for k in ("Source", "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal",
                                                   "L1 Wavefronts Shared Excessive", "Instructions Executed"))
        sm = H.index("# Samples")
        continue
    if H and len(r) > wi:
        try: rows.append((int(r[wi]), int(r[ii]), int(r[xi]), int(r[ei]), int(r[sm]), r[0][-5:], r[si].strip()))
        except ValueError: pass
tw = sum(x[0] for x in rows); ti = sum(x[1] for x in rows); te = sum(x[3] for x in rows); ts = sum(x[4] for x in rows)
print(f"{kern}: wavefronts {tw}, ideal {ti}, instr {te}, samples {ts}")
for w, i, x, e, s, a, src in sorted(rows, reverse=True)[:top]:
    print(f"{a} wf {100*w/tw:5.1f}%  ideal {100*i/tw:5.1f}%  wf/inst {w/max(e,1):5.2f}  samples {100*s/ts:4.1f}%  {src[:70]}")
