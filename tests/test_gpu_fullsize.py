"""Parity at the sizes the bench measures (VERDICT r1 "full-size parity"): the whole 8192 x 8192 bench DEM
(BASELINE.json configs[2], seed 20261104) through calc_form + calc_ddir2 + calc_shed4 on the device against the
oracle, every cell; calc_weti2, slope_gc and calc_relz + calc_length at 4096 x 4096 (the oracle needs ~1 min for
calc_weti2 at 8192^2).  Integers bit-exact, doubles bit-identical (NaN == NaN)."""
import os
import time

import numpy as np
import pytest

import litap_b200 as lb
import oracle
import relz_fixture as rf
from test_gpu_parity import FORM, WETI

pytestmark = pytest.mark.gpu

SEED = 20261104


def same_bits(a, b):
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    return ((a == b) | (np.isnan(a) & np.isnan(b))).all()


def test_bench_dem_8192_flow_and_form_vs_oracle():
    n = 8192
    t0 = time.time()
    g = lb.DeviceGrid(n + 2, n + 2)
    g.synth(seed=SEED)
    g.flow_form(5.0)                       # the call the bench times
    e = g.download("elev")
    # the device generator is the oracle's, bit for bit (rows checked here; the whole function in test_gpu_parity)
    ref_rows = oracle.synth_dem(64, n, seed=SEED, row0=4000, gnr=n)
    assert same_bits(e[4001:4065, :], ref_rows[1:-1, :])
    o_d = oracle.ddir(e)
    assert (g.download("ddir") == o_d["ddir"]).all(), "ddir"
    assert (g.download("flatcell") == o_d["flatcell"]).all(), "flatcell"
    o_s = oracle.shed(e, o_d["ddir"])
    assert g.npits == o_s["npits"]
    assert (g.download("shed") == o_s["shed"]).all(), "shed"
    assert (g.download("upslope").astype(np.int64) == o_s["upslope"]).all(), "upslope"
    assert (g.download("ridge") == o_s["ridge"]).all(), "ridge"
    del o_s, o_d
    o_f = oracle.form(e, 5.0, nthreads=os.cpu_count() or 1)
    for k in FORM:
        assert same_bits(g.download(k), o_f[k]), k
    g.close()
    print(f"8192^2 flow + form vs oracle: {time.time() - t0:.1f} s")


@pytest.fixture(scope="module")
def dem4096():
    n = 4096
    g = lb.DeviceGrid(n + 2, n + 2)
    g.synth(seed=SEED)
    g.ddir(); g.shed()
    out = {k: g.download(k) for k in ("elev", "ddir", "upslope", "shed")}
    out["grid"] = g
    yield out
    g.close()


def test_4096_weti_vs_oracle(dem4096):
    t0 = time.time()
    d = dem4096
    o = oracle.weti(d["elev"], d["ddir"], d["upslope"].astype(np.float64), 5.0)
    d["grid"].weti(5.0)
    flips = 0
    for k in WETI:
        a, b = d["grid"].download(k), o[k]
        assert (np.isnan(a) == np.isnan(b)).all(), f"{k}: NA pattern differs"
        bad = ~((a == b) | np.isnan(b))
        if bad.any():   # a 1-ulp libm log() difference may flip the truncated second decimal
            assert (np.abs(a[bad] - b[bad]) <= 0.01 * (1 + 1e-9)).all(), f"{k}: {bad.sum()} cells differ"
            flips += int(bad.sum())
    assert flips <= 16, f"{flips} truncation flips"
    print(f"4096^2 weti vs oracle: {time.time() - t0:.1f} s, {flips} flips, {lb.timings()['weti_ms']:.2f} ms on the device")


def test_4096_slope_gc_vs_oracle(dem4096):
    d = dem4096
    o = oracle.slope_gc(d["elev"], 5.0)
    d["grid"].slope_gc(5.0)
    for k in oracle.GC_F64:
        assert same_bits(d["grid"].download(k), o[k]), k
    for k in oracle.GC_U8 + oracle.GC_I32:
        assert (d["grid"].download(k) == o[k]).all(), k


def test_4096_relz_vs_oracle(dem4096):
    """calc_relz + calc_length on the flow fields of the 4096^2 DEM and of its inverse (host-buffer call)."""
    e = dem4096["elev"]
    NR, NC = e.shape

    def field(g, elev):
        d = g.download("ddir")
        seq = np.arange(1, NR * NC + 1, dtype=np.int64).reshape(NR, NC)
        drec = (seq + lb.api._OFF_DR[d] * NC + lb.api._OFF_DC[d]).astype(np.float64)
        drec[d == 0] = np.nan
        sh = g.download("shed").astype(np.float64)
        sh[sh == 0] = np.nan
        dd = d.astype(np.float64)
        dd[d == 0] = np.nan
        return {"elev": elev, "ddir": dd, "drec": drec, "upslope": g.download("upslope").astype(np.float64), "shed": sh}

    f = field(dem4096["grid"], e)
    gi = lb.DeviceGrid(NR, NC)
    gi.upload_elev(e)
    gi.ddir(invert=True); gi.shed()
    i = field(gi, gi.download("elev"))
    gi.close()
    db = dict(f, fill_shed=f["shed"], local_shed=f["shed"])
    idb = dict(i, inv_local_shed=i["shed"])
    got = lb.calc_relz(db, idb, 2000, 2000, pond=None, grid=5.0)
    want = oracle.relz(db, idb, None, 2000, 2000, 5.0)
    rf.assert_same(got, want, oracle.RELZ_NAMES, "4096^2")
