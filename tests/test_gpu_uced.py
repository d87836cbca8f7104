"""calc_upslopes(type = "uced") on the device against the oracle and the reference's goldens:
the three KATs of tests/testthat/test_02_flow_internal.R:83-153, flow/dem_fill.rds$uced (post pit removal: reversed
flow, merged sheds), synthetic DEMs incl. NA cells / inverted / plateaus.  Doubles must be bit-identical."""
import os

import numpy as np
import pytest

import litap_b200 as lb
import oracle
from test_oracle_golden import UCED_KATS

pytestmark = pytest.mark.gpu


def same_bits(a, b):
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    return ((a == b) | (np.isnan(a) & np.isnan(b))).all()


def device_uced(e, dd, sh):
    db = lb.add_buffer(e[1:-1, 1:-1])
    db["elev"] = np.ascontiguousarray(e).ravel()
    db["ddir_u8"] = np.ascontiguousarray(dd, dtype=np.uint8).ravel()
    s = sh.astype(np.float64).ravel()
    s[s == 0] = np.nan
    db["shedno"] = s
    return lb.calc_upslopes(db, type="uced")["uced"].reshape(e.shape)


@pytest.mark.parametrize("dem,uced", UCED_KATS)
def test_uced_kats(dem, uced):
    e = oracle.add_buffer(np.array(dem, dtype=float))
    d = oracle.ddir(e)["ddir"]
    s = oracle.shed(e, d)["shed"]
    got = device_uced(e, d, s)
    assert got[1:-1, 1:-1].ravel().tolist() == [float(x) for x in uced]
    assert np.isnan(got[0]).all()


def test_uced_dem_fill_golden(golden):
    g = golden["dem_fill"]
    e = oracle.add_buffer(g["elev"].reshape(90, 90))
    dd = np.nan_to_num(oracle.add_buffer(g["ddir"].reshape(90, 90)), nan=0).astype(np.uint8)
    sh = np.nan_to_num(oracle.add_buffer(g["fill_shed"].reshape(90, 90)), nan=0).astype(np.int32)
    got = device_uced(e, dd, sh)
    assert (got[1:-1, 1:-1].ravel() == g["uced"]).all()
    assert same_bits(got, oracle.uced(e, dd, sh))


@pytest.mark.parametrize("nr,nc,seed,quant,na,inv", [
    (300, 257, 3, 1000.0, 0, False),
    (129, 640, 5, 100.0, 0, False),     # 0.01 m: plateaus, zero drops
    (200, 200, 7, 1000.0, 10, False),   # 1 % NA cells
    (257, 300, 9, 1000.0, 0, True),     # inverted DEM
    (64, 64, 11, 0.0, 0, False),        # unquantised elevations: every difference rounds
])
def test_uced_synthetic_vs_oracle(nr, nc, seed, quant, na, inv):
    e = oracle.synth_dem(nr, nc, seed=seed, quant=quant, na_permille=na)
    if inv:
        e = oracle.invert(e)
    d = oracle.ddir(e)["ddir"]
    s = oracle.shed(e, d)["shed"]
    want = oracle.uced(e, d, s)
    assert same_bits(device_uced(e, d, s), want)


def test_uced_on_the_grid_and_in_ranges():
    """litap_grid_uced on a resident DEM (ddir / shed from the device), whole and with the contributions forced through
    several cell ranges (LITAP_UCED_PAIRS): same bits."""
    n = 1024
    g = lb.DeviceGrid(n + 2, n + 2)
    g.synth(seed=20261104)
    g.ddir(); g.shed()
    e, d, s = g.download("elev"), g.download("ddir"), g.download("shed")
    want = oracle.uced(e, d, s)
    g.uced()
    t = lb.timings()
    assert t["uced_ranges"] == 1 and t["uced_pairs"] > 0
    assert same_bits(g.download("uced"), want)
    os.environ["LITAP_UCED_PAIRS"] = str(int(t["uced_pairs"] // 5 + 1))
    try:
        g.uced()
        assert lb.timings()["uced_ranges"] >= 5
        assert same_bits(g.download("uced"), want)
    finally:
        del os.environ["LITAP_UCED_PAIRS"]
    g.close()


def test_uced_long_segments():
    """A long valley: a single flow line whose cells have no side inflow for > KSEG cells, so that segments longer than the
    per-thread buffer are re-read from the DEM."""
    nr, nc = 6, 200
    z = np.full((nr, nc), 500.0)
    z[2, :] = 400.0 - 0.37 * np.arange(nc)      # the channel, flowing east
    z[3, :] = 450.0 - 0.37 * np.arange(nc)      # one bank drains into it, the other rows are higher still
    e = oracle.add_buffer(z)
    d = oracle.ddir(e)["ddir"]
    s = oracle.shed(e, d)["shed"]
    assert same_bits(device_uced(e, d, s), oracle.uced(e, d, s))


def test_uced_errors():
    e = oracle.synth_dem(16, 16, seed=1)
    d = oracle.ddir(e)["ddir"]
    g = lb.DeviceGrid(18, 18)
    g.upload_elev(e)
    with pytest.raises(lb.LitapError):
        g.uced()          # no ddir / shed on the grid
    g.close()
    # a cycle in ddir must be reported, not walked for ever
    dd = d.copy()
    s = oracle.shed(e, d)["shed"]
    dd[5, 5], dd[5, 6] = 6, 4
    s[5, 5] = s[5, 6]
    with pytest.raises(lb.LitapError):
        device_uced(e, dd, s)
