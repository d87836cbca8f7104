import sys, os, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import litap_b200 as lb
n = 8192
g = lb.DeviceGrid(n + 2, n + 2); g.synth()
e = torch.empty((n + 2, n + 2), dtype=torch.float64).pin_memory(); e.numpy()[...] = g.download("elev"); g.close()
dts = {"ddir": torch.uint8, "flatcell": torch.uint8, "shed": torch.int32, "upslope": torch.int32, "ridge": torch.uint8,
       "slope_pct": torch.float64, "slope_deg": torch.float64, "aspect": torch.float64, "prof": torch.float64, "plan": torch.float64}
pinned = {k: torch.empty((n + 2, n + 2), dtype=dt).pin_memory() for k, dt in dts.items()}
out = {k: (v.numpy().view(np.uint32) if k == "upslope" else v.numpy()) for k, v in pinned.items()}
for i in range(4):
    t0 = time.perf_counter(); lb.flow_form(e.numpy(), 5.0, out=dict(out)); print("wall %.1f ms" % ((time.perf_counter() - t0) * 1e3), {k: round(v, 1) for k, v in lb.timings().items() if k in ("h2d_ms", "d2h_ms", "total_ms", "form_ms", "ddir_ms", "shed_ms")})
