#!/usr/bin/env python3
"""calc_weti2 on the bench DEM: stage time, dependency waves and per-kernel times.
usage: python profiles/weti_times.py [size] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import litap_b200 as lb  # noqa: E402
from litap_b200 import _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
_lib.check(lb.lib().litap_init(0))
g = lb.DeviceGrid(n + 2, n + 2)
g.synth()
g.ddir(); g.shed()
for it in range(reps):
    g.weti(5.0)
    t = lb.timings()
    print(f"rep {it}: weti {t['weti_ms']:.3f} ms  queued {t['weti_queued']}  launches {t['kernel_launches']}  "
          f"-> {n * n / t['weti_ms'] / 1e3:.1f} Mcells/s")
_lib.profile(True)
g.weti(5.0)
for k, (cnt, ms) in sorted(_lib.profile_report().items(), key=lambda x: -x[1][1]):
    print(f"  {k:20s} x{cnt:3d} {ms:9.3f} ms")
_lib.profile(False)
print("checksums", [hex(g.checksum(k)) for k in ("qarea1", "qarea2", "qweti1", "qweti2")])
g.close()
