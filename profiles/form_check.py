#!/usr/bin/env python3
"""Bit-for-bit check of the device calc_form against the CPU oracle on large synthetic DEMs
(test infrastructure: the oracle is the checker).  usage: python profiles/form_check.py [size]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import litap_b200 as lb  # noqa: E402
import oracle  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
lb._lib.check(lb.lib().litap_init(0))
nt = os.cpu_count() or 1
bad_total = 0
cases = [(20261104, 1000.0, 0, 5.0), (1, 1000.0, 0, 1.0), (2, 100.0, 0, 30.0), (3, 1000.0, 10, 5.0), (4, 1000.0, 0, 0.5),
         (5, 0.0, 0, 5.0)]
for seed, quant, na, grid in cases:
    e = oracle.synth_dem(n, n, seed=seed, quant=quant, na_permille=na, nthreads=nt)
    t0 = time.time()
    o = oracle.form(e, grid, nthreads=nt)
    t1 = time.time()
    g = lb.DeviceGrid(*e.shape)
    g.upload_elev(e)
    g.form(grid)
    ms = lb.timings()["form_ms"]
    bad = 0
    for k in ("slope_pct", "slope_deg", "aspect", "prof", "plan"):
        d = g.download(k)
        bad += int((d.view(np.uint64) != o[k].view(np.uint64)).sum() - (np.isnan(d) & np.isnan(o[k]) & (d.view(np.uint64) != o[k].view(np.uint64))).sum())
    g.close()
    bad_total += bad
    print(f"seed {seed} quant {quant} na {na} grid {grid}: {bad} differing values of {5 * e.size}; device {ms:.3f} ms, "
          f"oracle {t1 - t0:.1f} s ({nt} threads)", flush=True)
print("FORM_CHECK", "OK" if bad_total == 0 else f"FAILED ({bad_total})")
sys.exit(0 if bad_total == 0 else 1)
