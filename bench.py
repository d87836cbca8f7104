#!/usr/bin/env python3
"""bench.py — Mcells/s of the LITAP raster hot path (ddir + shed + upslope + form) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--size S] [--impl reference] [--weak]

A step = one pass of calc_ddir2 -> calc_shed4 (labels, upslope, ridge) + calc_form over one
synthetic hash-fBm DEM (0.001 m quantisation, grid = 5, seed 20261104) that is already resident in HBM.
  N = 1   BASELINE.json configs[2]: 8192 x 8192 on one GPU
  N > 1   BASELINE.json configs[3]: ONE 32768 x 32768 DEM cut into N row bands ("scaling": "strong"); after the
          timed region the bench proves the sharded bits: conservation (every cell drains into exactly one pit),
          npits equal on all ranks, and the band checksums of ddir / shed / upslope / ridge / slope_pct equal
          those of the same DEM processed unsharded on rank 0.  At N = 8 an `extra` block adds configs[4]:
          65536 x 65536 in 8 bands of 8192 rows.  --weak restores round 1's (N * 8192) x 8192 workload.
`value` = cells of the whole job / device time (CUDA events, max over ranks).  `e2e` = the same composite
through the host-buffer C-ABI call (pinned host arrays, H2D + D2H inside the timed region).
`--impl reference` times the CPU restatement of the reference's R code (the oracle; R itself is
not installed) on the N = 1 workload with all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED, QUANT, GRID_M = 20261104, 1000.0, 5.0
METRIC = "Mcells/s (ddir+shed+upslope+form)"
# algorithmic bytes per cell of each hot kernel
ALG_BYTES = {"k_d8": 9, "k_form": 48, "k_shed_local": 9, "k_shed_final": 10}
# roofline_frac_by_kernel charges a kernel with the helper launch that prepares its gathers
KERNEL_GROUP = {"k_shed_final": ("k_final_labels",)}


def ncu_traffic(kernel):
    """(GB, file) of 'traffic = dram read + write' for `kernel` in the latest (by name) profiles/ncu_*.txt (written by
    profiles/summarize.py from an ncu --set full capture of this bench at 8192 x 8192), else None."""
    import glob
    import re
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "ncu_*.txt")), reverse=True):  # names sort by round
        cur = None
        for line in open(f):
            if line.startswith("## "):
                cur = line
            elif cur and "traffic = dram read + write" in line and re.search(r"\b%s\(" % re.escape(kernel), cur):
                return float(line.split()[-2]) / 1e3, os.path.relpath(f, ROOT)
    return None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "20"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def wait_first(self, timeout=5.0):
        """nvidia-smi needs ~1 s before its first line: wait for it so that short timed regions are covered."""
        t0 = time.perf_counter()
        while self.p and not self.rows and time.perf_counter() - t0 < timeout:
            time.sleep(0.02)

    def stop(self, t_lo=None, t_hi=None):
        """Summary of the samples taken in [t_lo, t_hi] (perf_counter times; the loaded region)."""
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        self.t.join(timeout=2)
        allrows = self.rows
        self.rows = [r for t, r in allrows if (t_lo is None or t >= t_lo) and (t_hi is None or t <= t_hi)]
        if not self.rows:  # region shorter than the sampling period: nearest samples
            self.rows = [r for _, r in allrows[-3:]]
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_port_pass(e, nthreads):
    """One pass of the oracle (CPU port of the reference's R code) over a buffered DEM."""
    import oracle
    d = oracle.ddir(e)
    oracle.shed(e, d["ddir"])
    oracle.form(e, GRID_M, nthreads=nthreads)


def cpu_sample(size, rows):
    import numpy as np

    import oracle
    rows = min(rows, size)
    e = oracle.synth_dem(rows, size, seed=SEED, quant=QUANT, row0=0, gnr=size)
    e[-1, :] = np.nan  # the band is processed as a DEM of its own: close the ring
    what = (f"the whole {size}x{size} DEM" if rows == size else
            f"rows 0-{rows - 1} of the same {size}x{size} DEM") + f" ({rows * size / 1e6:.1f} Mcells)"
    return e, rows * size, what


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nthreads = os.cpu_count() or 1
    if a.gpus > 1 and not a.weak and not a.size:
        # N > 1: our arm runs the 32768^2 DEM; the CPU arm takes a bounded sample of THAT DEM with as many cells as the
        # whole 8192^2 workload (its first 2048 rows: ~5 s per pass with 16 threads)
        a.size = a.dem
        rows = a.cpu_rows or 2048
    else:
        # N = 1 (or --weak): the whole 8192^2 workload unless --cpu-rows bounds the sample (12-13 Mcells/s with 16 threads)
        a.size = a.size or 8192
        rows = a.cpu_rows or a.size
    e, cells, what = cpu_sample(a.size, rows)
    for _ in range(min(a.warmup, 1)):
        cpu_port_pass(e, nthreads)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        cpu_port_pass(e, nthreads)
    dt = (time.perf_counter() - t0) / a.steps
    v = cells / dt / 1e6
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mcells/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": min(a.warmup, 1), "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong" if (a.gpus > 1 and not a.weak and a.size == a.dem) else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"hash-fBm DEM {a.size}x{a.size}, quant 0.001 m, grid 5 m, seed {SEED}",
                   "note": "R is not installed: CPU port of the reference's R algorithm (oracle/), "
                           "OpenMP where the algorithm is data-parallel (D8 stencil, form), serial elsewhere"},
        "cpu_baseline": {"value": v, "unit": "Mcells/s", "cores": nthreads, "kind": "port", "sample": what},
        "e2e": {"value": v, "unit": "Mcells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def run_ours(a):
    # libraries (NCCL's version banner, ...) may write to stdout: keep fd 1 for the ONE JSON line
    real_out = os.dup(1)
    os.dup2(2, 1)
    import numpy as np
    import torch
    import torch.distributed as dist

    import litap_b200 as lb
    from litap_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.check(lb.lib().litap_init(local))
    stream = torch.cuda.current_stream()
    _lib.set_stream(stream.cuda_stream)

    if world > 1:
        # ONE DEM cut into row bands, one band per GPU; halo rows and boundary tables travel over NCCL inside the
        # library (litap_b200/csrc/bands.cu).  Default: BASELINE.json configs[3], 32768 x 32768 (strong scaling);
        # --weak: (world * size) x size, 8192-row bands of 8192 columns as in round 1.
        from litap_b200.bands import BandedDEM, band_rows, init_comm
        init_comm(rank, world)
        if a.weak or a.size:
            n = a.size or 8192                    # rows per band
            ncols = a.cols or n
            gnr = n * world
        else:
            gnr = ncols = a.dem
            n = gnr // world
        scaling = "weak" if (a.weak or a.size) else "strong"
        g = BandedDEM(gnr, ncols, world, band=rank)
        g.synth(seed=SEED, quant=QUANT)
    else:
        n = a.size or 8192
        ncols = a.cols or n
        gnr = n
        scaling = "weak"
        g = lb.DeviceGrid(n + 2, ncols + 2)
        g.synth(seed=SEED, quant=QUANT)
    NR, NC = n + 2, ncols + 2

    fused = not a.serial_stages

    def step():
        if fused:
            g.flow_form(GRID_M)  # one call: calc_form overlaps the flow stages on a second stream
        else:
            g.form(GRID_M)
            g.ddir()
            g.shed()

    def step_serial():  # the per-kernel profiling pass times every kernel alone on the device
        g.form(GRID_M)
        g.ddir()
        g.shed()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.wait_first()
    t_load0 = time.perf_counter()
    for _ in range(a.warmup):
        step()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    nl = 0
    ev0.record(stream)
    for _ in range(a.steps):
        if fused:
            g.flow_form(GRID_M); nl += lb.timings()["kernel_launches"]
        else:
            g.form(GRID_M); nl += lb.timings()["kernel_launches"]
            g.ddir(); nl += lb.timings()["kernel_launches"]
            g.shed(); nl += lb.timings()["kernel_launches"]
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop(t_load0, time.perf_counter()) if sampler else None
    if clocks:
        clocks["window"] = "warm-up + timed steps"
    t = torch.tensor([ms], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    cells_job = gnr * ncols
    value = cells_job * a.steps / (ms * 1e-3) / 1e6
    stage = lb.timings()

    # ---- the sharded bits (outside the timed region): conservation, npits, band checksums
    CHECKED = ("ddir", "shed", "upslope", "ridge", "slope_pct")
    checks = None
    band_sums = None
    if world > 1:
        cells_b, pits_b = g.conservation()
        mine = [cells_b, pits_b, int(g.npits)] + [g.band_checksum(k) for k in CHECKED]
        allv = [None] * world
        dist.all_gather_object(allv, mine)
        cells_all, pits_all = sum(v[0] for v in allv), sum(v[1] for v in allv)
        band_sums = {k: sum(v[3 + i] for v in allv) & 0xFFFFFFFFFFFFFFFF for i, k in enumerate(CHECKED)}
        checks = {"cells_with_a_shed": cells_all, "sum_upslope_at_pits": pits_all,
                  "conservation": cells_all == pits_all and cells_all == cells_job,
                  "npits": allv[0][2], "npits_equal_on_all_ranks": len({v[2] for v in allv}) == 1}

    # ---- per-kernel pass (untimed, CUDA events around every launch) -> roofline of the dominant kernel
    _lib.profile(True)
    PROF_STEPS = 3
    for _ in range(PROF_STEPS):
        step_serial()
    prof = _lib.profile_report()
    _lib.profile(False)
    per_kernel = {k: {"launches_per_step": v[0] / PROF_STEPS, "ms_per_step": v[1] / PROF_STEPS} for k, v in prof.items()}
    dom = max(prof, key=lambda k: prof[k][1])
    peak, peak_src = measured_peak()
    roof = {"bound": "hbm", "kernel": dom, "achieved": None, "peak": peak, "unit": "GB/s", "frac": None,
            "traffic": None, "peak_source": peak_src}
    if dom in ALG_BYTES:
        avg_ms = prof[dom][1] / prof[dom][0]
        ach = ALG_BYTES[dom] * n * ncols / (avg_ms * 1e-3) / 1e9
        roof.update(achieved=ach, frac=ach / peak, alg_bytes_per_cell=ALG_BYTES[dom], avg_launch_ms=avg_ms)
    # DRAM bytes of one launch of that kernel from the committed ncu --set full capture (same workload), if any
    if world == 1 and n == 8192 and ncols == 8192:
        tr = ncu_traffic(dom)
        if tr:
            roof.update(traffic=tr[0], traffic_unit="GB per launch (dram read + write)", traffic_source=tr[1])
    rf_all = {}
    for k, b in ALG_BYTES.items():
        if k in prof:
            avg = prof[k][1] / prof[k][0]
            for extra in KERNEL_GROUP.get(k, ()):   # helper launches that belong to the kernel's job
                if extra in prof:
                    avg += prof[extra][1] / prof[k][0]
            rf_all[k] = round(b * n * ncols / (avg * 1e-3) / 1e9 / peak, 4)

    # ---- calc_weti2 (not part of the composite metric; single-device stage), reported beside it
    weti = None
    if world == 1 and not a.no_weti:
        g.weti(GRID_M)
        g.weti(GRID_M)
        tw = lb.timings()
        weti = {"ms": tw["weti_ms"], "Mcells_per_s": n * n / tw["weti_ms"] / 1e3, "cells_through_ready_queues": tw["weti_queued"],
                "alg_bytes_per_cell": 45, "frac_of_hbm_peak": round(45 * n * n / (tw["weti_ms"] * 1e-3) / 1e9 / peak, 4),
                "note": "order-dependent fp64 accumulation, bit-exact vs the reference order; barrier-free dataflow, latency-bound"}

    # ---- end to end through the host-buffer C ABI (pinned host arrays)
    e2e = None
    e2e_note = None
    if not a.no_e2e and world > 1:
        import psutil
        need = (n + 3) * NC * 8 + (n + 1) * NC * 51
        if need * world > 0.5 * psutil.virtual_memory().available:
            e2e_note = f"skipped: {need * world / 1e9:.0f} GB of pinned host buffers exceed half of the host's free memory"
    if not a.no_e2e and world > 1 and e2e_note is None:
        # every rank: its band's rows from pinned host memory -> stages -> its rows of every output back
        hr = torch.empty((n + 3, NC), dtype=torch.float64).pin_memory()
        _lib.check(lb.lib().litap_grid_download(g._h[0], b"elev", hr.numpy().ctypes.data))
        rows_in = hr.numpy()[1:]
        _, nown = g.rows(0)
        dts = {"ddir": torch.uint8, "flatcell": torch.uint8, "shed": torch.int32, "upslope": torch.int32,
               "ridge": torch.uint8, "slope_pct": torch.float64, "slope_deg": torch.float64,
               "aspect": torch.float64, "prof": torch.float64, "plan": torch.float64}
        pinned = {k: torch.empty((nown, NC), dtype=dt).pin_memory() for k, dt in dts.items()}

        def e2e_step():
            g.upload_rows(rows_in)
            step()
            for k, v in pinned.items():
                g.download_into(k, v.numpy())

        h2d = rows_in.nbytes * world
        d2h = sum(v.numel() * v.element_size() for v in pinned.values()) * world
        e2e_step()
        barrier()
        es = max(1, min(a.steps, a.e2e_steps if n * ncols <= 2 ** 27 else 2))
        ev0.record(stream)
        for _ in range(es):
            e2e_step()
        ev1.record(stream)
        barrier()
        t = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item()) / es
        e2e = {"value": cells_job / (e2e_ms * 1e-3) / 1e6, "unit": "Mcells/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "steps": es,
               "api": "litap_band_upload_rows + litap_bands_form/ddir/shed + litap_band_download (C ABI, pinned host buffers, all ranks)"}
        del pinned, hr
    elif not a.no_e2e:
        host_in = torch.empty((NR, NC), dtype=torch.float64).pin_memory()
        e_np = host_in.numpy()
        e_np[...] = g.download("elev")
        dts = {"ddir": torch.uint8, "flatcell": torch.uint8, "shed": torch.int32, "upslope": torch.int32,
               "ridge": torch.uint8, "slope_pct": torch.float64, "slope_deg": torch.float64,
               "aspect": torch.float64, "prof": torch.float64, "plan": torch.float64}
        pinned = {k: torch.empty((NR, NC), dtype=dt).pin_memory() for k, dt in dts.items()}
        out = {k: (v.numpy().view(np.uint32) if k == "upslope" else v.numpy()) for k, v in pinned.items()}
        h2d = e_np.nbytes
        d2h = sum(v.nbytes for v in out.values())
        for _ in range(1):
            lb.flow_form(e_np, GRID_M, out=out)
        barrier()
        es = max(1, min(a.steps, a.e2e_steps))
        ev0.record(stream)
        for _ in range(es):
            lb.flow_form(e_np, GRID_M, out=out)
        ev1.record(stream)
        barrier()
        t = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item()) / es
        e2e = {"value": cells_job / (e2e_ms * 1e-3) / 1e6, "unit": "Mcells/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "steps": es,
               "api": "litap_flow_form (C ABI, pinned host buffers)",
               "wire_d2h_bytes_per_step": (d2h - 20 * NR * NC) if os.environ.get("LITAP_FORM_PACK", "1") != "0" else d2h,
               "note": "d2h_bytes_per_step = bytes delivered into the caller's arrays; the five calc_form outputs cross PCIe "
                       "as int32 thousandths and are widened to fp64 by host threads (bit-identical), unless LITAP_FORM_PACK=0"}
        del pinned
        # the same call with PAGEABLE buffers, fresh result arrays every call: what an R caller (fresh R vectors) sees
        e_page = np.array(e_np)
        del host_in, e_np
        os.environ.setdefault("LITAP_D2H_HUGEPAGE", "1")   # opt-in: huge-page advice on the fresh result arrays
        pw = []
        for _ in range(3):
            t0 = time.perf_counter()
            res = lb.flow_form(e_page, GRID_M)
            pw.append((time.perf_counter() - t0) * 1e3)
            del res
        e2e["pageable"] = {"ms_per_step": min(pw[1:]), "value": cells_job / (min(pw[1:]) * 1e-3) / 1e6, "unit": "Mcells/s",
                           "what": "litap_flow_form with numpy (pageable) input and freshly allocated numpy outputs, "
                                   "wall clock, best of 2 after 1 warm-up; staged through pinned buffers by the library; "
                                   "LITAP_D2H_HUGEPAGE=" + os.environ.get("LITAP_D2H_HUGEPAGE", "0")}
        del e_page

    # ---- file to files: .flt on disk -> HBM -> composite -> ten raw binary files on disk
    files = None
    if world == 1 and not a.no_e2e and not a.no_files:
        import shutil
        import tempfile
        base = None
        for cand in ("/dev/shm", None):
            try:
                d = tempfile.mkdtemp(prefix="litap_bench_", dir=cand)
                if shutil.disk_usage(d).free > 60 * n * ncols + (1 << 30):
                    base = d
                    break
                shutil.rmtree(d, ignore_errors=True)
            except OSError:
                pass
        if base:
            try:
                g.write_raw("elev", os.path.join(base, "dem.flt"), as_flt=True)
                outs = ("ddir", "flatcell", "shed", "upslope", "ridge", "slope_pct", "slope_deg", "aspect", "prof", "plan")
                best = None
                for _ in range(2):
                    t0 = time.perf_counter()
                    gf = lb.DeviceGrid.read_flt(os.path.join(base, "dem.flt"))
                    t1 = time.perf_counter()
                    gf.flow_form(GRID_M)
                    t2 = time.perf_counter()
                    gf.write_raw_many({k: os.path.join(base, k + ".bin") for k in outs})
                    t3 = time.perf_counter()
                    gf.close()
                    cur = {"read_ms": (t1 - t0) * 1e3, "compute_ms": (t2 - t1) * 1e3, "write_ms": (t3 - t2) * 1e3,
                           "wall_ms": (t3 - t0) * 1e3}
                    if best is None or cur["wall_ms"] < best["wall_ms"]:
                        best = cur
                nbytes = sum(os.path.getsize(os.path.join(base, k + ".bin")) for k in outs)
                files = dict(best, value=cells_job / (best["wall_ms"] * 1e-3) / 1e6, unit="Mcells/s",
                             bytes_read=os.path.getsize(os.path.join(base, "dem.flt")), bytes_written=nbytes,
                             where=("/dev/shm (RAM-backed: the file system is not the limit)" if base.startswith("/dev/shm")
                                    else "the box's temporary directory"),
                             api="litap_grid_read_flt + litap_grid_flow_form + litap_grid_write_raw_many (10 files, 8 writers), best of 2")
                lb.lib().litap_pool_trim()
            finally:
                shutil.rmtree(base, ignore_errors=True)

    # ---- slope_gc (not part of the composite metric), reported beside it
    gc = None
    if world == 1 and not a.no_weti:
        g.slope_gc(GRID_M)
        g.slope_gc(GRID_M)
        tg = lb.timings()
        # compulsory traffic: elev read once, 6 f64 + 2 u8 + 4 i32 outputs written once = 74 B/cell
        gc = {"ms": tg["gc_ms"], "Mcells_per_s": n * n / tg["gc_ms"] / 1e3, "alg_bytes_per_cell": 74,
              "frac_of_hbm_peak": round(74 * n * n / (tg["gc_ms"] * 1e-3) / 1e9 / peak, 4)}

    # ---- calc_upslopes(type = "uced") (not part of the composite metric)
    uced = None
    if world == 1 and not a.no_weti:
        g.uced()
        g.uced()
        tu = lb.timings()
        uced = {"ms": tu["uced_ms"], "Mcells_per_s": n * n / tu["uced_ms"] / 1e3, "alg_bytes_per_cell": 21,
                "frac_of_hbm_peak": round(21 * n * n / (tu["uced_ms"] * 1e-3) / 1e9 / peak, 4),
                "contributions": tu["uced_pairs"], "ranges": tu["uced_ranges"],
                "note": "bit-exact restatement of the reference's per-track long double sums: work = the sum over all cells of "
                        "the cells upslope of them, not one pass over the DEM; alg. bytes = elev 8 + ddir 1 + shed 4 + uced 8"}
        lb.lib().litap_pool_trim()

    # ---- pit_stat1 (not part of the composite metric): pour points of the DEM's initial sheds
    pitstat = None
    if world == 1 and not a.no_weti:
        g.pit_stat("landmapr")
        tbl = g.pit_stat("landmapr")
        tp = lb.timings()
        pitstat = {"ms": tp["pitstat_ms"], "Mcells_per_s": n * n / tp["pitstat_ms"] / 1e3, "sheds": int(len(tbl["shedno"])),
                   "method": "landmapr", "note": "device part of pit_stat1 on the resident grid (table of 21 columns back to the host included)"}
        del tbl
        lb.lib().litap_pool_trim()

    # ---- calc_relz + calc_length (not part of the composite metric) on the flow field of a
    # 4096^2 DEM through the host-buffer C ABI (11 input grids up, 39 double grids down)
    relz = None
    if world == 1 and not a.no_weti:
        m = min(n, 4096)
        rdb, ridb = relz_flow_fields(m)
        for _ in range(2):
            t0 = time.perf_counter()
            out = lb.calc_relz(rdb, ridb, 1000, 1000, pond=None, grid=GRID_M)
            wall = (time.perf_counter() - t0) * 1e3
        tz = lb.timings()
        relz = {"workload": f"flow field of the hash-fBm DEM {m}x{m}, str_val = ridge_val = 1000", "device_ms": tz["relz_ms"],
                "Mcells_per_s_device": (m + 2) ** 2 / tz["relz_ms"] / 1e3, "h2d_ms": tz["h2d_ms"], "d2h_ms": tz["d2h_ms"],
                "call_wall_ms": wall, "jump_rounds": tz["plateau_rounds"], "kernel_launches": tz["kernel_launches"],
                "longest_path": int(np.nanmax(out["n2pit"])), "host_buffers": "pageable numpy arrays"}
        del rdb, ridb, out

    # ---- CPU baseline (rank 0, N = 1): the oracle on a bounded sample, single thread
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu:
        e, cells, what = cpu_sample(n, a.cpu_rows or 2048)
        t0 = time.perf_counter()
        cpu_port_pass(e, 1)
        dt = time.perf_counter() - t0
        nth = os.cpu_count() or 1
        t0 = time.perf_counter()
        cpu_port_pass(e, nth)
        dta = time.perf_counter() - t0
        cpu = {"value": cells / dt / 1e6, "unit": "Mcells/s", "cores": 1, "kind": "port", "sample": what,
               "seconds": dt,
               "all_cores": {"value": cells / dta / 1e6, "cores": nth, "seconds": dta,
                             "note": "OpenMP in the data-parallel stages (D8 stencil, calc_form); plateau rounds, patches and calc_shed4 serial"}}

    # ---- N > 1: the same DEM unsharded on rank 0 (it fits one B200 up to 32768^2): checksums must equal the bands'
    extra = None
    if world > 1:
        g.close()
        lb.lib().litap_pool_trim()
        fits = (gnr + 2) * (((ncols + 2 + 31 + 31) // 32) * 32) < 2 ** 31 and gnr + 2 <= 65535 and gnr * ncols * 130 < 170e9
        if rank == 0 and fits and not a.no_whole:
            t0 = time.perf_counter()
            w = lb.DeviceGrid(gnr + 2, ncols + 2)
            w.synth(seed=SEED, quant=QUANT)
            w.flow_form(GRID_M)
            whole = {k: w.checksum(k) for k in CHECKED}
            wn = int(w.npits)
            w.close()
            lb.lib().litap_pool_trim()
            checks["sharded_equals_whole"] = all(whole[k] == band_sums[k] for k in CHECKED) and wn == checks["npits"]
            checks["whole_dem"] = {"what": f"the same {gnr}x{ncols} DEM as one grid on rank 0 (litap_grid_flow_form), "
                                           f"checksums of {', '.join(CHECKED)} and npits", "seconds": time.perf_counter() - t0,
                                   "checksums": {k: hex(v) for k, v in whole.items()},
                                   "band_checksums": {k: hex(v) for k, v in band_sums.items()}}
        elif rank == 0:
            checks["sharded_equals_whole"] = None
            checks["whole_dem"] = "not run: the DEM does not fit one device" if not fits else "not run (--no-whole)"
        barrier()
        # ---- N = 8: BASELINE.json configs[4], 65536 x 65536 in 8 bands of 8192 rows (device-resident, generated on the device)
        if world == 8 and scaling == "strong" and not a.no_64k:
            g64 = BandedDEM(65536, 65536, world, band=rank)
            g64.synth(seed=SEED, quant=QUANT)
            for _ in range(2):
                g64.flow_form(GRID_M)
            barrier()
            k64 = max(2, min(a.steps, 5))
            ev0.record(stream)
            for _ in range(k64):
                g64.flow_form(GRID_M)
            ev1.record(stream)
            barrier()
            t = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms64 = float(t.item()) / k64
            c64 = g64.conservation()
            allv = [None] * world
            dist.all_gather_object(allv, [c64[0], c64[1], int(g64.npits)])
            cells64, pits64 = sum(v[0] for v in allv), sum(v[1] for v in allv)
            extra = {"workload": "hash-fBm DEM 65536x65536 in 8 row bands of 8192 rows (BASELINE.json configs[4])",
                     "ms_per_step": ms64, "steps": k64, "value": 65536 * 65536 / (ms64 * 1e-3) / 1e6, "unit": "Mcells/s",
                     "npits": allv[0][2], "npits_equal_on_all_ranks": len({v[2] for v in allv}) == 1,
                     "cells_with_a_shed": cells64, "sum_upslope_at_pits": pits64,
                     "conservation": cells64 == pits64 == 65536 * 65536,
                     "note": "upslope is a 32-bit count: litap_bands_shed refuses a DEM in which a basin reaches 2^32 cells "
                             "(this conservation test, inside the library)"}
            # end to end at this size: every band's rows up from pinned host memory, the step, every result down.  The ten
            # results of a band are 31.7 GB; they go through ONE pinned buffer per rank (the largest result, 4.3 GB) in
            # turn, so that the host needs 8 x 8.6 GB of pinned memory instead of 8 x 36 GB: same bytes over PCIe.
            if not a.no_e2e:
                import psutil
                r0, nown = g64.rows(0)
                NC64 = 65536 + 2
                need = (8192 + 3) * NC64 * 8 + nown * NC64 * 8
                if need * world * 2 < psutil.virtual_memory().available:
                    hin = torch.empty((8192 + 3, NC64), dtype=torch.float64).pin_memory()
                    _lib.check(lb.lib().litap_grid_download(g64._h[0], b"elev", hin.numpy().ctypes.data))
                    rows64 = hin.numpy()[1:]
                    hout = torch.empty((nown, NC64), dtype=torch.float64).pin_memory()
                    views = {"ddir": np.uint8, "flatcell": np.uint8, "shed": np.int32, "upslope": np.uint32, "ridge": np.uint8,
                             "slope_pct": np.float64, "slope_deg": np.float64, "aspect": np.float64, "prof": np.float64,
                             "plan": np.float64}
                    flat = hout.numpy().reshape(-1).view(np.uint8)

                    def e2e64():
                        g64.upload_rows(rows64)
                        g64.flow_form(GRID_M)
                        for k, dt in views.items():
                            nb = nown * NC64 * np.dtype(dt).itemsize
                            g64.download_into(k, flat[:nb].view(dt).reshape(nown, NC64))

                    e2e64()
                    barrier()
                    ev0.record(stream)
                    e2e64()
                    ev1.record(stream)
                    barrier()
                    t = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    up_b = rows64.nbytes * world
                    dn_b = sum(nown * NC64 * np.dtype(dt).itemsize for dt in views.values()) * world
                    extra["e2e"] = {"ms_per_step": float(t.item()), "value": 65536 * 65536 / (float(t.item()) * 1e-3) / 1e6,
                                    "unit": "Mcells/s", "h2d_bytes_per_step": up_b, "d2h_bytes_per_step": dn_b, "steps": 1,
                                    "GBps_total": (up_b + dn_b) / (float(t.item()) * 1e-3) / 1e9,
                                    "note": "per-band pinned host buffers; the ten results of a band pass through one pinned "
                                            "buffer in turn (host memory), same bytes over PCIe"}
                    del hin, hout
                else:
                    extra["e2e"] = {"value": None, "note": "skipped: not enough free host memory for the pinned buffers"}
            g64.close()

    if rank == 0:
        if world == 1:
            wl = f"hash-fBm DEM {n}x{ncols} (BASELINE.json configs[2])"
        elif scaling == "strong":
            wl = f"hash-fBm DEM {gnr}x{ncols} in {world} row bands of {n} rows (BASELINE.json configs[3])"
        else:
            wl = f"hash-fBm DEM {gnr}x{ncols} in {world} row bands of {n} rows (--weak)"
        os.write(real_out, (json.dumps({
            "metric": METRIC, "value": value, "unit": "Mcells/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl + f", quant 0.001 m, grid 5 m, seed {SEED}",
                       "cells_per_gpu": n * ncols,
                       "l2": f"inputs ({n * ncols * 8 / 1e6:.0f} MB elev per GPU) larger than L2; no flush needed",
                       "parallelism": "1 GPU" if world == 1 else
                       f"row bands, one per GPU; NCCL halo rows + boundary-graph tables (DESIGN.md section 4)",
                       "step": ("litap_grid_flow_form / litap_bands_flow_form: calc_form on the caller's stream beside calc_ddir2 + calc_shed4 on a "
                                "high-priority stream; per_kernel / roofline come from a separate serial pass"
                                if fused else "calc_form, calc_ddir2, calc_shed4 as three calls"),
                       "npits": stage["npits"], "flat_cells": stage["flat_cells"]},
            "clocks": clocks, "e2e": e2e if e2e is not None else ({"value": None, "note": e2e_note} if e2e_note else None),
            "gpu_launches": nl, "checks": checks, "extra": extra,
            "roofline": roof, "roofline_frac_by_kernel": rf_all, "per_kernel": per_kernel, "file_to_files": files, "weti": weti, "slope_gc": gc, "uced": uced, "pit_stat": pitstat, "relz": relz,
            "cpu_baseline": cpu}) + "\n").encode())
    if world == 1:
        g.close()
    if world > 1:
        from litap_b200.bands import destroy_comm
        destroy_comm()
        dist.destroy_process_group()


def relz_flow_fields(n):
    """db / idb for calc_relz from the device stages on an n x n synthetic DEM: elev, ddir, drec, upslope and the
    initial sheds (used as fill_shed / local_shed / inv_local_shed) of the DEM and of the inverted DEM."""
    import numpy as np
    import litap_b200 as lb
    from litap_b200.api import _OFF_DC, _OFF_DR
    g = lb.DeviceGrid(n + 2, n + 2)
    g.synth(SEED)
    out = []
    sh = (n + 2, n + 2)
    for inv in (False, True):
        g.ddir(invert=inv)          # invert = True turns elev over in place first
        g.shed()
        d = g.download("ddir").reshape(sh)
        seq = np.arange(1, d.size + 1, dtype=np.int64).reshape(sh)
        drec = (seq + _OFF_DR[d].astype(np.int64) * (n + 2) + _OFF_DC[d]).astype(np.int32)
        drec[d == 0] = 0
        out.append({"elev": g.download("elev").reshape(sh), "ddir": d, "drec": drec,
                    "upslope": g.download("upslope").reshape(sh).astype(np.int32), "shed": g.download("shed").reshape(sh)})
    g.close()
    f, i = out
    return dict(f, fill_shed=f["shed"], local_shed=f["shed"]), dict(i, inv_local_shed=i["shed"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--size", type=int, default=0, help="DEM rows per GPU (and columns unless --cols); default 8192 at N = 1; "
                                                        "at N > 1 it selects the weak-scaling workload")
    ap.add_argument("--cols", type=int, default=0, help="DEM columns (multi-GPU band mode), default = --size")
    ap.add_argument("--dem", type=int, default=32768, help="N > 1: side of the one DEM that is cut into N row bands")
    ap.add_argument("--weak", action="store_true", help="N > 1: (N * size) x size cells, one size-row band per GPU")
    ap.add_argument("--no-whole", action="store_true", help="N > 1: skip the unsharded run of the same DEM on rank 0")
    ap.add_argument("--no-64k", action="store_true", help="N = 8: skip the 65536 x 65536 extra block")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-rows", type=int, default=0, help="rows of the DEM in the CPU sample (default: 2048 for the in-line cpu_baseline, the whole DEM for --impl reference)")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-files", action="store_true", help="skip the file-to-files measurement")
    ap.add_argument("--no-weti", action="store_true")
    ap.add_argument("--serial-stages", action="store_true", help="form, ddir, shed as three calls (no stream overlap)")
    a = ap.parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
