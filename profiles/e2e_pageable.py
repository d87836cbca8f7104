#!/usr/bin/env python3
"""litap_flow_form through the C ABI with PAGEABLE host buffers (what an R caller has), 8192^2:
the driver's own pageable copy path (LITAP_STAGED_D2H=0) against the staged one (hostcopy.cu).
usage: [LITAP_STAGED_D2H=0] python profiles/e2e_pageable.py [size]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import litap_b200 as lb  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
lb._lib.check(lb.lib().litap_init(0))
g = lb.DeviceGrid(n + 2, n + 2)
g.synth()
e = g.download("elev").reshape(n + 2, n + 2)
g.close()
for it in range(3):
    t0 = time.perf_counter()
    out = lb.flow_form(e, 5.0)          # fresh numpy outputs every call, like fresh R vectors
    wall = (time.perf_counter() - t0) * 1e3
    t = lb.timings()
    print(f"staged={os.environ.get('LITAP_STAGED_D2H', '1')} run {it}: wall {wall:.0f} ms = {n * n / wall / 1e3:.0f} Mcells/s "
          f"(h2d {t['h2d_ms']:.0f} ms, d2h tail {t['d2h_ms']:.0f} ms, total {t['total_ms']:.0f} ms)")
    del out
