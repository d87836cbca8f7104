#!/usr/bin/env python3
"""Device time (CUDA events around the stage) vs host wall time of every stage call on the
bench workload: shows where a step spends time outside kernels (syncs, host work).
usage: python profiles/stage_times.py [size] [steps]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import litap_b200 as lb  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
lb._lib.check(lb.lib().litap_init(0))
g = lb.DeviceGrid(n + 2, n + 2)
g.synth()
for it in range(steps):
    row = []
    for name, fn, key in (("form", lambda: g.form(5.0), "form_ms"), ("ddir", g.ddir, "ddir_ms"),
                          ("shed", g.shed, "shed_ms")):
        t0 = time.perf_counter()
        fn()
        wall = (time.perf_counter() - t0) * 1e3
        t = lb.timings()
        row.append(f"{name}: dev {t[key]:.3f} wall {wall:.3f} launches {t['kernel_launches']}")
    t = lb.timings()
    print(f"step {it}: " + " | ".join(row) + f" | rounds {t['plateau_rounds']} flats {t['flat_cells']} "
          f"patch {t['patch_cells']} npits {t['npits']}")
g.close()
