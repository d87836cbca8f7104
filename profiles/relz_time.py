#!/usr/bin/env python3
"""litap_relz (calc_relz + calc_length) on the flow field of the bench DEM: device time, copies, per-kernel
profile, and the oracle on a smaller grid for scale.
usage: python profiles/relz_time.py [size=4096] [oracle_size=1024]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import litap_b200 as lb  # noqa: E402
from bench import relz_flow_fields as flow_fields  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
m = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
lb._lib.check(lb.lib().litap_init(0))
db, idb = flow_fields(n)
lb._lib.profile(True)
for it in range(3):
    t0 = time.perf_counter()
    out = lb.calc_relz(db, idb, 1000, 1000, pond=None, grid=5.0)
    wall = (time.perf_counter() - t0) * 1e3
    t = lb.timings()
    cells = (n + 2) ** 2
    print(f"relz {n}^2 run {it}: device {t['relz_ms']:.2f} ms ({cells / t['relz_ms'] / 1e3:.0f} Mcells/s), "
          f"h2d {t['h2d_ms']:.1f} ms, d2h {t['d2h_ms']:.1f} ms, wall {wall:.0f} ms, jump rounds {t['plateau_rounds']}, "
          f"launches {t['kernel_launches']}, longest path {int(np.nanmax(out['n2pit']))}")
print(lb._lib.profile_report())
if m:
    db, idb = flow_fields(m)
    import oracle
    t0 = time.perf_counter()
    want = oracle.relz(db, idb, None, 1000, 1000, 5.0)
    dt = time.perf_counter() - t0
    got = lb.calc_relz(db, idb, 1000, 1000, pond=None, grid=5.0)
    same = all(np.array_equal(got[k], want[k], equal_nan=True) for k in oracle.RELZ_NAMES)
    print(f"oracle {m}^2: {dt * 1e3:.0f} ms ({(m + 2) ** 2 / dt / 1e6:.2f} Mcells/s, 1 thread); identical to the device: {same}")
