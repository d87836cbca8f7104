#!/usr/bin/env python3
"""calc_upslopes(type = "uced") on the bench DEM: stage time, contributions and per-kernel times.
usage: python profiles/uced_times.py [size] [reps]"""
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
    g.uced()
    t = lb.timings()
    print(f"rep {it}: uced {t['uced_ms']:.3f} ms  contributions {t['uced_pairs']:.0f} ({t['uced_pairs'] / n / n:.2f} per cell) "
          f"ranges {t['uced_ranges']} -> {n * n / t['uced_ms'] / 1e3:.1f} Mcells/s")
_lib.profile(True)
g.uced()
for k, (cnt, ms) in sorted(_lib.profile_report().items(), key=lambda x: -x[1][1]):
    print(f"  {k:20s} x{cnt:3d} {ms:9.3f} ms")
_lib.profile(False)
print("checksum", hex(g.checksum("uced")))
g.close()
