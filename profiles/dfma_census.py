#!/usr/bin/env python3
"""Per-kernel census of fused multiply-adds in the shipped library (the build uses -fmad=false because the
reference never contracts a*b+c; the DFMAs that remain must all come from IEEE division / sqrt expansions (they sit
next to a MUFU.RCP64H / MUFU.RSQ64H seed) or from explicit fma() calls in cellmath.h).
usage: python profiles/dfma_census.py > profiles/dfma_census_r2.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "litap_b200", "liblitap_cuda.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
cur, stats = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)::", "").split("(")[0]
        stats.setdefault(cur, collections.Counter())
        continue
    if cur is None:
        continue
    for op in ("DFMA", "DMUL", "DADD", "MUFU.RCP64H", "MUFU.RSQ64H", "FFMA", "UBLKCP", "UTMALDG"):
        if re.search(r"\b%s\b" % re.escape(op), line):
            stats[cur][op] += 1
# explicit fma() / __fma_rn sites in the sources, by file
explicit = {}
for f in sorted(os.listdir(os.path.join(ROOT, "litap_b200", "csrc"))):
    if f.endswith((".cu", ".cuh", ".h")):
        txt = open(os.path.join(ROOT, "litap_b200", "csrc", f)).read()
        n = len(re.findall(r"\bfma\s*\(|__fma_rn\s*\(", txt))
        if n:
            explicit[f] = n
print("# DFMA census of litap_b200/liblitap_cuda.so (cuobjdump -sass; build flags: -fmad=false)")
print("# explicit fma() call sites in the sources:", explicit)
print("# a kernel whose DFMAs all belong to division / sqrt expansions has MUFU.RCP64H / RSQ64H seeds; kernels that call")
print("# div_rcp / div1000 / the quick paths of cellmath.h (explicit fma()) are marked in the last column")
FMA_USERS = {"k_form": "form_cell_quick / fast (explicit fma, guarded; the doubt queue falls back to the reference tree)",
             "k_form_edges": "same as k_form", "k_d8": "div_sqrt2 (exact quotient of the largest diagonal difference: 4 explicit fma per cell x 4 cells) + the IEEE divisions of d8_slow_at",
             "d8_slow_at": "IEEE divisions of the reference loop (out-of-line slow path of k_d8)", "k_weti_nodes": "div_rcp", "k_weti_flow": "div_rcp", "k_weti_cls": "div_rcp",
             "k_weti_final": "log() of libdevice", "k_relz_cells": "r_round1 (div1000-style exact quotient)"}
print(f"{'kernel':44s} {'DFMA':>6s} {'DMUL':>6s} {'DADD':>6s} {'RCP64H':>7s} {'RSQ64H':>7s} {'UBLKCP':>7s}  explicit fma()")
for k, c in stats.items():
    if not (c["DFMA"] or c["DMUL"] or c["DADD"]):
        continue
    short = k.split("::")[-1]
    short = re.sub(r"<.*", "", short).replace("void ", "")
    print(f"{short[:44]:44s} {c['DFMA']:6d} {c['DMUL']:6d} {c['DADD']:6d} {c['MUFU.RCP64H']:7d} {c['MUFU.RSQ64H']:7d} {c['UBLKCP']:7d}  "
          f"{FMA_USERS.get(short, '')}")
tot = sum(c["DFMA"] for c in stats.values())
print(f"# total DFMA {tot}; kernels with DFMA but neither a division / sqrt seed nor a listed explicit-fma path:",
      [k.split('::')[-1] for k, c in stats.items() if c["DFMA"] and not c["MUFU.RCP64H"] and not c["MUFU.RSQ64H"]
       and re.sub(r"<.*", "", k.split("::")[-1]).replace("void ", "") not in FMA_USERS])
