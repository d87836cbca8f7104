#!/usr/bin/env python3
"""Per-kernel device times (CUDA events around every launch) of the composite step on the bench DEM, plus the
checksums of its integer outputs, for one build of the library (LITAP_SO selects a tuning build).
usage: [LITAP_SO=...] python profiles/kernel_times.py [size] [steps] [kernel-substring ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import litap_b200 as lb  # noqa: E402
from litap_b200 import _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
want = sys.argv[3:]
_lib.check(lb.lib().litap_init(0))
g = lb.DeviceGrid(n + 2, n + 2)
g.synth()
for _ in range(2):
    g.form(5.0); g.ddir(); g.shed()
_lib.profile(True)
for _ in range(steps):
    g.form(5.0); g.ddir(); g.shed()
prof = _lib.profile_report()
_lib.profile(False)
tot = sum(v[1] for v in prof.values()) / steps
out = {k: round(v[1] / steps, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1]) if not want or any(w in k for w in want)}
sums = {k: g.checksum(k) for k in ("ddir", "flatcell", "shed", "upslope", "ridge", "slope_pct")}
print(os.environ.get("LITAP_SO", "default"), "sum_ms", round(tot, 4), out)
print("  checksums", sums, "npits", int(g.npits))
g.close()
